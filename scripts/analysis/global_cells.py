"""Offline sizing of the round-2 layout (model in banded_model.py
add_global_cells): with the direct edges of every row folded into "global cells", how many values
does phase 2 stream per edge, and how many bytes does each phase move?  R-MAT from the numpy
oracle, degree-renumbered, kMinCell = 3, hot-band threshold 32 K edges, cells cut every 2048
stream positions.  usage: python scripts/analysis/global_cells.py [scale] [band_entries]"""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from oracle import graph_oracle as go  # noqa: E402

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
V, E = 1 << scale, 16 << scale
t0 = time.time()
src, dst = go.rmat_edgelist(scale, E, seed=0)
indeg = np.bincount(dst, minlength=V)
order = np.argsort(-indeg, kind="stable")
newid = np.empty(V, np.int64)
newid[order] = np.arange(V)
key = (newid[dst] << 32) | newid[src]
del src, dst
key.sort()
row = (key >> 32).astype(np.int32)
col = (key & 0xFFFFFFFF).astype(np.int32)
del key
print(f"scale {scale}: V {V}, E {E}, built in {time.time() - t0:.0f} s", flush=True)

band = col // B
f = np.ones(E, bool)
f[1:] = (row[1:] != row[:-1]) | (band[1:] != band[:-1])
f |= (np.arange(E) % 2048) == 0
starts = np.flatnonzero(f)
sizes = np.diff(np.append(starts, E))
cb = band[starts]
big = sizes >= 3
per_band = np.bincount(cb[big], weights=sizes[big], minlength=int(band.max()) + 1)
hot = per_band >= 32768
pre = big & hot[cb]                      # pre-reduced cells
n_banded = int(sizes[pre].sum())
n_bcells = int(pre.sum())
direct_edge = np.repeat(~pre, sizes)
drow = row[direct_edge]
n_direct = int(drow.size)
gstart = np.ones(n_direct, bool)
gstart[1:] = drow[1:] != drow[:-1]
gstart |= (np.arange(n_direct) % 2048) == 0
n_gcells = int(gstart.sum())
rows_nz = int((indeg > 0).sum())
vb = 4  # float32
print(f"band {B} entries: pre-reduced edges {n_banded / E:.3f} E in {n_bcells / E:.4f} E cells; "
      f"direct edges {n_direct / E:.3f} E in {n_gcells / E:.4f} E global cells "
      f"({n_gcells / max(rows_nz, 1):.2f} per non-empty row)")
now = (E - n_banded + n_bcells) / E
print(f"phase 2 today: {now:.3f} E entries (index + gather each); with global cells: "
      f"{(n_bcells + n_gcells) / E:.3f} E partials streamed, no indices")
p1_bytes = n_banded * 2 + n_bcells * (4 + vb) + n_direct * 4 + n_gcells * (4 + vb)
p2_bytes = (n_bcells + n_gcells) * vb + rows_nz * (4 + vb + vb)
csc_bytes = E * 4 + V * (4 + 2 * vb)
print(f"bytes per SpMV (f32, unweighted): phase 1 {p1_bytes / 1e6:.0f} MB "
      f"(16-bit band stream + 32-bit global stream + partial stores), phase 2 {p2_bytes / 1e6:.0f} MB; "
      f"the plain CSC the roofline counts: {csc_bytes / 1e6:.0f} MB")
