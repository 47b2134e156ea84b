"""Offline: share of pre-reducible edges and remaining phase-2 gathers as a function of the band
size (entries of x per band), kMinCell = 3, hot-band threshold 32 K edges.  Decides whether larger
f64 bands (28 K entries = 224 KB instead of 16 K) are worth a non-power-of-two band."""
import sys
import numpy as np
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 24
s = np.load(f"/tmp/s{scale}_s.npy"); d = np.load(f"/tmp/s{scale}_d.npy")
E = s.size
key = (d.astype(np.int64) << 32) | s.astype(np.int64)
key.sort()
row = (key >> 32).astype(np.int32); src = (key & 0xffffffff).astype(np.int32); del key
for B in (8192, 16384, 24576, 28672, 32768, 49152, 65536):
    band = src // B
    f = np.ones(E, bool)
    f[1:] = (row[1:] != row[:-1]) | (band[1:] != band[:-1])
    starts = np.flatnonzero(f)
    sizes = np.diff(np.append(starts, E))
    cb = band[starts]
    m = sizes >= 3
    per_band = np.bincount(cb[m], weights=sizes[m], minlength=int(band.max()) + 1)
    hot = per_band >= 32768
    mm = m & hot[cb]
    eb = sizes[mm].sum(); pc = int(mm.sum())
    print(f"band {B:6d} entries: pre-reduced {eb / E:.3f}  partials {pc / E:.4f}  phase-2 entries "
          f"{(E - eb + pc) / E:.3f}  hot bands {int(hot.sum())}", flush=True)
