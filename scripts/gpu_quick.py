"""ms per PageRank iteration at R-MAT scale 24 for (weighted, dtype) combinations with the
library selected by B200_LIBRARY / the environment's knobs.  Usage: gpu_quick.py [scale] [tag]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cugraph_forked_b200 import engine  # noqa: E402

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 24
tag = sys.argv[2] if len(sys.argv) > 2 else ""
combos = sys.argv[3].split(",") if len(sys.argv) > 3 else ["u32", "u64", "w32", "w64"]
h = engine.Handle()
nv, ne = 1 << scale, 16 << scale
s, d = engine.generate_rmat_edgelist(h, scale, ne, seed=0)
verts = torch.arange(nv, dtype=torch.int32, device="cuda")
for c in combos:
    weighted = c[0] == "w"
    dt = torch.float32 if c[1:] == "32" else torch.float64
    w = torch.rand(ne, device="cuda", dtype=torch.float32) if weighted else None
    g = engine.Graph(h, s, d, w, vertices=verts)
    engine.pagerank(g, dtype=dt, fixed_iterations=5)
    _, st = engine.pagerank(g, dtype=dt, fixed_iterations=40)
    print(f"{tag} {c}: {st.elapsed_ms / 40:.4f} ms/iter  {ne * 40 / st.elapsed_ms / 1e6:.1f} GTEPS "
          f"launches/iter={st.gpu_launches / 40:.0f}", flush=True)
    del g
