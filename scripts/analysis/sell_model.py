"""Offline sizing of the round-2 layout ("lane owns a segment"): phase 1 = one lane per band cell
(cells sorted by length inside a band, 32 cells per slice, indices interleaved), phase 2 = one lane
per row of the reduced CSC (32 consecutive rows per slice).  Prints entries and padding for band
sizes / minimum cell sizes.  Input: /tmp/s{scale}_{s,d}.npy from gen_edges.py."""
import sys
import numpy as np
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
s = np.load(f"/tmp/s{scale}_s.npy"); d = np.load(f"/tmp/s{scale}_d.npy")
E = s.size; V = 1 << scale
key = (d.astype(np.int64) << 32) | s.astype(np.int64)
key.sort()
row = (key >> 32).astype(np.int32); src = (key & 0xffffffff).astype(np.int32); del key
def slice_pad(lens, gran=1, sort_window=0):
    """entries stored when `lens` are packed 32 per slice, each padded to the slice maximum"""
    l = ((lens + gran - 1) // gran) * gran
    n = (len(l) + 31) // 32 * 32
    l = np.concatenate([l, np.zeros(n - len(l), l.dtype)])
    if sort_window:
        w = (len(l) + sort_window - 1) // sort_window * sort_window
        l = np.concatenate([l, np.zeros(w - len(l), l.dtype)]).reshape(-1, sort_window)
        l = -np.sort(-l, axis=1); l = l.reshape(-1)
    m = l.reshape(-1, 32).max(axis=1)
    return int(m.sum()) * 32
for B, minc, cellmax in [(32768, 3, 2048), (32768, 2, 2048), (49152, 2, 2048), (57344, 2, 2048), (16384, 3, 2048), (16384, 2, 2048), (28672, 2, 2048)]:
    band = src // B
    pos = np.arange(E)
    f = np.ones(E, bool)
    f[1:] = (row[1:] != row[:-1]) | (band[1:] != band[:-1])
    f |= (pos % cellmax) == 0
    starts = np.flatnonzero(f)
    sizes = np.diff(np.append(starts, E))
    cb = band[starts]; crow = row[starts]
    m = sizes >= minc
    per_band = np.bincount(cb[m], weights=sizes[m], minlength=int(band.max()) + 1)
    hot = per_band >= 32768
    mm = m & hot[cb]
    eb = int(sizes[mm].sum()); pc = int(mm.sum())
    # phase 1: cells sorted by length inside their band
    order = np.lexsort((-sizes[mm], cb[mm]))
    cl = sizes[mm][order]; cbb = cb[mm][order]
    # slices never cross bands: pad per band
    p1 = {}
    for gran in (1, 2, 4):
        tot = 0
        bstart = np.flatnonzero(np.r_[True, cbb[1:] != cbb[:-1]])
        bend = np.r_[bstart[1:], len(cbb)]
        for a, b in zip(bstart, bend):
            tot += slice_pad(cl[a:b], gran)
        p1[gran] = tot
    # phase 2: reduced row lengths = direct edges + partials per row
    direct_per_row = np.bincount(crow[~mm], weights=sizes[~mm], minlength=V).astype(np.int64)
    part_per_row = np.bincount(crow[mm], minlength=V).astype(np.int64)
    rl = direct_per_row + part_per_row
    nz = int((np.bincount(row, minlength=V) > 0).sum())
    rl_nz = rl[:nz]
    long_thr = 128
    reg = np.where(rl_nz > long_thr, 0, rl_nz)
    nlong = int((rl_nz > long_thr).sum()); long_entries = int(rl_nz[rl_nz > long_thr].sum())
    print(f"B={B} minc={minc}: banded {eb/E:.3f}E cells {pc/E:.4f}E | phase1 stored idx: gran1 {p1[1]/E:.3f}E gran2 {p1[2]/E:.3f}E gran4 {p1[4]/E:.3f}E"
          f" | reduced entries {(E-eb+pc)/E:.3f}E direct {(E-eb)/E:.3f}E; long rows(>{long_thr}) {nlong} with {long_entries/E:.4f}E;"
          f" phase2 stored: consecutive {slice_pad(reg)/E:.3f}E sorted1024 {slice_pad(reg,1,1024)/E:.3f}E sorted4096 {slice_pad(reg,1,4096)/E:.3f}E", flush=True)
