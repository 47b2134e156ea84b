"""HITS on R-MAT: device time per iteration (two pull SpMVs: authorities over the CSC, hubs over
the CSR companion) and the HBM-roofline fraction of one SpMV.  usage: gpu_hits_bench.py <scale> [f32|f64]"""
import json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cugraph_forked_b200 import engine
scale = int(sys.argv[1]); dt = sys.argv[2] if len(sys.argv) > 2 else "f32"
tdt = torch.float32 if dt == "f32" else torch.float64
nv, ne = 1 << scale, 16 << scale
h = engine.Handle()
s, d = engine.generate_rmat_edgelist(h, scale, ne, seed=0)
g = engine.Graph(h, s, d, vertices=torch.arange(nv, dtype=torch.int32, device="cuda"))
engine.hits(g, dtype=tdt, fixed_iterations=3)
iters = 20
_, _, st = engine.hits(g, dtype=tdt, fixed_iterations=iters)
ms = st.elapsed_ms / iters
vs = 4 if dt == "f32" else 8
b = (nv + 1) * 4 + ne * 4 + ne * vs
peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
print(json.dumps({"algo": "hits", "scale": scale, "dtype": dt, "ms_per_iteration": ms,
                  "gteps": 2 * ne / ms / 1e6, "spmv_roofline_frac": 2 * b / (ms * 1e-3) / 1e9 / peak,
                  "note": "an iteration = 2 SpMVs + max-normalisations + L1 difference (host reads one scalar)"}))
