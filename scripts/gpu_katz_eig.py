"""Katz / eigenvector centrality on R-MAT scale 24 (same SpMV, other epilogues): ms per iteration."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cugraph_forked_b200 import engine
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 24
h = engine.Handle(); nv, ne = 1 << scale, 16 << scale
s, d = engine.generate_rmat_edgelist(h, scale, ne)
g = engine.Graph(h, s, d, vertices=torch.arange(nv, dtype=torch.int32, device="cuda")); del s, d
for dt, name in ((torch.float32, "f32"), (torch.float64, "f64")):
    for _ in range(2):
        _, st = engine.katz_centrality(g, alpha=1e-7, beta=1.0, dtype=dt, fixed_iterations=20)
    print(f"katz s{scale} {name}: {st.elapsed_ms/20:.4f} ms/iter {ne/(st.elapsed_ms/20)/1e6:.1f} GTEPS", flush=True)
    for _ in range(2):
        _, st = engine.eigenvector_centrality(g, dtype=dt, fixed_iterations=20)
    print(f"eigenvector s{scale} {name}: {st.elapsed_ms/20:.4f} ms/iter {ne/(st.elapsed_ms/20)/1e6:.1f} GTEPS", flush=True)
