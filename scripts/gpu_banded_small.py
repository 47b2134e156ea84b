"""Small banded run for compute-sanitizer: R-MAT scale 15, raw SpMV (f32/f64, weighted and not)
and a few PageRank iterations with B200_BANDED=1."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("B200_BANDED", "1")
from cugraph_forked_b200 import engine  # noqa: E402

h = engine.Handle()
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 15
nv, ne = 1 << scale, 16 << scale
s, d = engine.generate_rmat_edgelist(h, scale, ne, seed=1)
verts = torch.arange(nv, dtype=torch.int32, device="cuda")
for weighted in (False, True):
    w = torch.rand(ne, device="cuda", dtype=torch.float32) if weighted else None
    g = engine.Graph(h, s, d, w, vertices=verts)
    for dt in (torch.float32, torch.float64):
        x = torch.rand(nv, device="cuda", dtype=dt)
        y = engine.spmv_pull(g, x)
        offs = g.offsets().long()
        idx = g.indices().long()
        rows = torch.repeat_interleave(torch.arange(nv, device="cuda"), offs[1:] - offs[:-1])
        vals = x.double()[idx] * (g.weights().double() if weighted else 1.0)
        ref = torch.zeros(nv, device="cuda", dtype=torch.float64).index_add_(0, rows, vals)
        err = (torch.abs(y.double() - ref) / ref.abs().clamp_min(1e-30)).max().item()
        pr, st = engine.pagerank(g, dtype=dt, fixed_iterations=3)
        print(f"weighted={weighted} dtype={dt} spmv_err={err:.3e} launches={st.gpu_launches} "
              f"sum={pr.sum().item():.9f}")
print("SMALL_OK")
