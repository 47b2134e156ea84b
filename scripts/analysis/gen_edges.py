"""Offline input for the layout analyses: R-MAT edge list (the generator's own stream, chunked),
renumbered by descending in-degree (ties by id) like the graph build, saved as /tmp/s{scale}_{s,d}.npy."""
import sys
import numpy as np
sys.path.insert(0, ".")
from oracle import graph_oracle
scale = int(sys.argv[1]); E = 16 << scale; V = 1 << scale
chunk = 1 << 24
ss, dd = [], []
for f in range(0, E, chunk):
    s, d = graph_oracle.rmat_edgelist(scale, min(chunk, E - f), seed=0, first_edge=f)
    ss.append(s); dd.append(d)
s = np.concatenate(ss); d = np.concatenate(dd); del ss, dd
deg = np.bincount(d, minlength=V)
order = np.lexsort((np.arange(V), -deg))     # descending in-degree, ties by id
new = np.empty(V, np.int32); new[order] = np.arange(V, dtype=np.int32)
np.save(f"/tmp/s{scale}_s.npy", new[s]); np.save(f"/tmp/s{scale}_d.npy", new[d])
print("done", scale, E)
