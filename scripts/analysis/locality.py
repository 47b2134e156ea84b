"""Offline locality analysis of the R-MAT gather stream (no GPU): how many 32-byte sectors of x
the pull kernel must fetch per edge under different edge orders / cache sizes.  Uses the numpy
oracle's generator and renumbering (test infrastructure)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from oracle import graph_oracle as go

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
ef = 16
V, E = 1 << scale, ef << scale
t0 = time.time()
src, dst = go.rmat_edgelist(scale, E, seed=0)
print("rmat", time.time() - t0, flush=True)
indeg = np.bincount(dst, minlength=V)
order = np.argsort(-indeg, kind="stable")          # new id -> old id
newid = np.empty(V, np.int64); newid[order] = np.arange(V)
s = newid[src].astype(np.int32); d = newid[dst].astype(np.int32)
del src, dst
deg = indeg[order]
print("V", V, "E", E, "max indeg", deg[0])
for thr in (1024, 32, 1):
    rows = int((deg >= thr).sum()); print(f"rows deg>={thr}: {rows}  edges {deg[:rows].sum()/E:.3f}")
# source coverage: fraction of edges with src < H
cs = np.cumsum(np.bincount(s, minlength=V)) / E
for H in (16<<10, 32<<10, 48<<10, 64<<10, 128<<10, 256<<10, 1<<20, 2<<20, 4<<20, 7<<20):
    if H <= V: print(f"src < {H:8d} ({H/V:6.2%} of V): {cs[H-1]:.3f} of edges")
np.save(f"/tmp/s{scale}_s.npy", s); np.save(f"/tmp/s{scale}_d.npy", d)
