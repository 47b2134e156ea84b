"""Sweep of the banded path's knobs on one R-MAT graph (built once per layout knob):
phase-2 tile configuration (B200_PULL_CFG), hub-cache budget (B200_SMEM_BUDGET_KB), hot-band
threshold (B200_BAND_HOT_MIN).  Prints ms per PageRank iteration."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cugraph_forked_b200 import engine  # noqa: E402

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 24
dtype = torch.float32 if (len(sys.argv) < 3 or sys.argv[2] == "f32") else torch.float64
h = engine.Handle()
nv, ne = 1 << scale, 16 << scale
s, d = engine.generate_rmat_edgelist(h, scale, ne, seed=0)
verts = torch.arange(nv, dtype=torch.int32, device="cuda")


def run(g, iters=40):
    engine.pagerank(g, dtype=dtype, fixed_iterations=5)
    _, st = engine.pagerank(g, dtype=dtype, fixed_iterations=iters)
    return st.elapsed_ms / iters, st.spmv_ms / max(st.spmv_launches, 1)


def setenv(**kw):
    for k, v in kw.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = str(v)


setenv(B200_BANDED=0)
g = engine.Graph(h, s, d, vertices=verts)
print("tiled           : %.4f ms/iter (spmv %.4f)" % run(g), flush=True)
setenv(B200_BANDED=1)
for hot in (32768, 8192, 131072, 524288):
    setenv(B200_BAND_HOT_MIN=hot, B200_PULL_CFG=None, B200_SMEM_BUDGET_KB=None)
    g = engine.Graph(h, s, d, vertices=verts)
    print("banded hot_min=%7d : %.4f ms/iter (spmv %.4f)" % ((hot,) + run(g)), flush=True)
setenv(B200_BAND_HOT_MIN=None)
g = engine.Graph(h, s, d, vertices=verts)
run(g)
if dtype == torch.float32:
    for cfg in (0, 1, 2, 3, 4, 5):
        setenv(B200_PULL_CFG=cfg)
        print("banded cfg=%d      : %.4f ms/iter (spmv %.4f)" % ((cfg,) + run(g)), flush=True)
    setenv(B200_PULL_CFG=None)
for kb in (110, 130, 160, 190, 210):
    setenv(B200_SMEM_BUDGET_KB=kb)
    print("banded smem=%3d KB : %.4f ms/iter (spmv %.4f)" % ((kb,) + run(g)), flush=True)
setenv(B200_SMEM_BUDGET_KB=None)
for il in (0, 1):
    setenv(B200_TILE_INTERLEAVE=il)
    g2 = engine.Graph(h, s, d, vertices=verts)
    print("banded interleave=%d: %.4f ms/iter (spmv %.4f)" % ((il,) + run(g2)), flush=True)
print("SWEEP_OK")
