"""Development aid: time fixed-iteration PageRank on an RMAT graph (device-resident inputs)."""
import sys, os, time, argparse
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cugraph_forked_b200 import engine

ap = argparse.ArgumentParser()
ap.add_argument("--scale", type=int, default=24)
ap.add_argument("--ef", type=int, default=16)
ap.add_argument("--iters", type=int, default=20)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--dtype", default="f32")
ap.add_argument("--weighted", action="store_true")
a = ap.parse_args()
h = engine.Handle()
ne = a.ef << a.scale
dt = torch.float32 if a.dtype == "f32" else torch.float64
t0 = time.time()
src, dst = engine.generate_rmat_edgelist(h, a.scale, ne)
w = engine.uniform_fill(h, ne, seed=1, dtype=dt) if a.weighted else None
verts = torch.arange(1 << a.scale, dtype=torch.int32, device="cuda")
torch.cuda.synchronize(); t1 = time.time()
g = engine.Graph(h, src, dst, w, vertices=verts)
torch.cuda.synchronize(); t2 = time.time()
del src, dst
print(f"s{a.scale} gen {t1-t0:.3f}s build {t2-t1:.3f}s seg={g.segment_offsets()}", flush=True)
V = 1 << a.scale
sz = 4 if a.dtype == "f32" else 8
bytes_iter = (V + 1) * 4 + ne * 4 + ne * sz + (ne * sz if a.weighted else 0)
for r in range(a.reps):
    pr, st = engine.pagerank(g, dtype=dt, fixed_iterations=a.iters)
    ms = st.elapsed_ms / a.iters
    print(f"  rep{r}: {ms:.4f} ms/iter  {ne/ms/1e6:.1f} GTEPS  algorithmic {bytes_iter/ms/1e6:.1f} GB/s "
          f"({bytes_iter/ms/1e6/6539.5*100:.1f}% of measured 6539.5)", flush=True)

# tuning builds compiled with -DB200_PROFILE_ROLES export per-CTA role clocks of the last launch
import ctypes
from cugraph_forked_b200 import _lib
if hasattr(_lib.lib(), "b200_debug_role_clocks"):
    import numpy as np
    buf = (ctypes.c_ulonglong * (256 * 8))()
    _lib.lib().b200_debug_role_clocks(buf)
    a = np.frombuffer(buf, dtype=np.uint64).reshape(256, 8)[:148].astype(np.float64)
    names = ["kernel", "gather wait full", "gather zero rows", "reduce wait full",
             "reduce wait ready", "producer wait empty", "tiles", "-"]
    for k, nme in enumerate(names[:7]):
        print(f"  role clocks {nme:22s} mean {a[:, k].mean():12.0f}  min {a[:, k].min():12.0f}  "
              f"max {a[:, k].max():12.0f}")
