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
g = engine.Graph(h, s, d, vertices=verts)
run(g)
if dtype == torch.float32:
    for hub in (None, 0):
        for cfg in range(11):
            setenv(B200_PULL_CFG=cfg, B200_HUB_ENTRIES=hub)
            print("banded cfg=%2d hub=%-4s : %.4f ms/iter (spmv %.4f)" % ((cfg, hub) + run(g)), flush=True)
    setenv(B200_PULL_CFG=None, B200_HUB_ENTRIES=None)
else:
    for hub in (None, 0, 2048):
        setenv(B200_HUB_ENTRIES=hub)
        print("banded hub=%-4s : %.4f ms/iter (spmv %.4f)" % ((hub,) + run(g)), flush=True)
    setenv(B200_HUB_ENTRIES=None)
print("SWEEP_OK")
