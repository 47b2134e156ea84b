"""Small end-to-end run for compute-sanitizer memcheck: sell layout build + both phases (f32/f64,
weighted/unweighted), HITS (companions), the no-alloc SpMV; scale 14."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from cugraph_forked_b200 import engine
h = engine.Handle()
scale = 14; nv, ne = 1 << scale, 16 << scale
s, d = engine.generate_rmat_edgelist(h, scale, ne, seed=1)
w = torch.rand(ne, device="cuda")
for weights in (None, w):
    g = engine.Graph(h, s, d, weights, vertices=torch.arange(nv, dtype=torch.int32, device="cuda"))
    for dt in (torch.float32, torch.float64):
        pr, st = engine.pagerank(g, dtype=dt, fixed_iterations=5)
        assert abs(pr.double().sum().item() - 1.0) < 1e-4
        engine.katz_centrality(g, alpha=1e-4, dtype=dt, max_iterations=50)
    hubs, auth, _ = engine.hits(g, fixed_iterations=3)
    x = torch.rand(nv, device="cuda")
    y = engine.spmv_pull(g, x, 1.0, 0.0)
print("SANITY_OK")
