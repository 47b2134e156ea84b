"""Offline model of the banded pre-reduction (no GPU): for a degree-renumbered R-MAT graph, cut
every row's ascending neighbour list into cells = runs of sources inside one band of B entries
of x; cells with >= MINC edges are pre-reduced from the band's shared-memory copy (phase 1),
the rest are gathered directly (phase 2).  Prints how many gathers phase 2 has left."""
import sys
import numpy as np
scale = int(sys.argv[1]); B = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
s = np.load(f"/tmp/s{scale}_s.npy"); d = np.load(f"/tmp/s{scale}_d.npy")
E = s.size; V = 1 << scale
key = (d.astype(np.int64) << 32) | s.astype(np.int64)
key.sort()
row = (key >> 32).astype(np.int32); src = (key & 0xffffffff).astype(np.int32); del key
band = src // B
f = np.ones(E, bool)
f[1:] = (row[1:] != row[:-1]) | (band[1:] != band[:-1])
starts = np.flatnonzero(f)
sizes = np.diff(np.append(starts, E))
cband = band[starts]
print(f"E={E} cells={starts.size} ({starts.size/E:.3f}/edge)")
for minc in (2, 3, 4, 6, 8, 16):
    m = sizes >= minc
    eb = sizes[m].sum(); pc = int(m.sum())
    print(f"MINC={minc:2d}: banded edges {eb/E:.3f}  partials {pc/E:.4f}  phase-2 gathers {(E-eb+pc)/E:.3f}  "
          f"edges/partial {eb/max(pc,1):.1f}")
minc = 3
m = sizes >= minc
nb = (V + B - 1) // B
eb_band = np.bincount(cband[m], weights=sizes[m], minlength=nb)
pc_band = np.bincount(cband[m], minlength=nb)
cum = np.cumsum(eb_band) / E
for k in (0, 1, 2, 3, 7, 15, 31, 63, 127, 255, nb - 1):
    if k < nb: print(f"band {k:4d}: edges {eb_band[k]/E:.4f} cum {cum[k]:.3f} partials {pc_band[k]}")
print("bands with any banded edge:", int((eb_band > 0).sum()), "of", nb)
