"""Executable model (numpy, no GPU) of the round-2 "lane owns a segment" layout and of both
kernels' arithmetic (csrc/sell.cu, csrc/sell.cuh): the same cell / slice / slot / window rules,
small sizes only.  tests/test_sell_model.py checks it against a plain CSC SpMV and checks the
layout's invariants; the CUDA build follows these rules step by step.

  build(offsets, indices, weights, ...)  -> layout dict
  spmv(layout, x)                        -> y = A^T-pull of x through phase 1 and phase 2
"""
import numpy as np

G = 4            # 16-bit indices per lane and load in phase 1 (kSellG)


def build(offsets, indices, weights=None, *, band_entries=64, min_cell=2, cell_max=8, hot_min=1,
          window=64, long_rows=6, skip_empty=False):
    offsets = np.asarray(offsets, np.int64)
    indices = np.asarray(indices, np.int64)
    ne, nr = len(indices), len(offsets) - 1
    rows = np.repeat(np.arange(nr), np.diff(offsets))
    band = indices // band_entries
    # 1. cell starts: first edge of a row, band change, or a position that is a multiple of cell_max
    start = np.ones(ne, bool)
    if ne > 1:
        start[1:] = (rows[1:] != rows[:-1]) | (band[1:] != band[:-1]) | (np.arange(1, ne) % cell_max == 0)
    cell_first = np.flatnonzero(start)
    cell_len = np.diff(np.append(cell_first, ne))
    cell_band = band[cell_first]
    # 2. candidates (>= min_cell edges) of hot bands (>= hot_min candidate edges) are pre-reduced
    cand = cell_len >= min_cell
    nb = int(band.max()) + 1 if ne else 0
    per_band = np.bincount(cell_band[cand], weights=cell_len[cand], minlength=nb)
    pre = cand & (per_band[cell_band] >= hot_min) if ne else cand
    # 3. partial slots: row-major rank q of the pre-reduced cells; reduced CSC = direct edges + one
    #    entry per pre-reduced cell (index x_split + q), in CSC order
    q_of_cell = np.cumsum(pre) - 1
    n_part = int(pre.sum())
    x_split = int(indices.max()) + 1 if ne else 0
    red_idx, red_w, red_row = [], [], []
    for c, (f, l) in enumerate(zip(cell_first, cell_len)):
        if pre[c]:
            red_idx.append(x_split + q_of_cell[c]); red_w.append(1.0); red_row.append(rows[f])
        else:
            for e in range(f, f + l):
                red_idx.append(indices[e]); red_w.append(1.0 if weights is None else weights[e])
                red_row.append(rows[e])
    red_idx = np.array(red_idx, np.int64); red_w = np.array(red_w, np.float64)
    red_row = np.array(red_row, np.int64)
    roff = np.zeros(nr + 1, np.int64)
    np.cumsum(np.bincount(red_row, minlength=nr), out=roff[1:])
    # 4. phase 1: cells sorted by (band, length descending), stable; 32 per slice, never across
    #    bands; slots follow the SORTED order (band-major); indices interleaved G at a time
    pc = np.flatnonzero(pre)
    order = pc[np.lexsort((np.arange(len(pc)), -cell_len[pc], cell_band[pc]))]
    inv = np.empty(n_part, np.int64)          # row-major q -> band-major slot
    p1 = []                                   # slices: dict(band, cells=[(first, len, slot)], groups)
    slot = 0
    for b in np.unique(cell_band[order]) if len(order) else []:
        cs = order[cell_band[order] == b]
        for s0 in range(0, len(cs), 32):
            sl = cs[s0:s0 + 32]
            cells = []
            for c in sl:
                inv[q_of_cell[c]] = slot
                cells.append((int(cell_first[c]), int(cell_len[c]), slot))
                slot += 1
            p1.append(dict(band=int(b), cells=cells, groups=-(-int(cell_len[sl[0]]) // G)))
    part_ref = red_idx >= x_split
    red_idx[part_ref] = x_split + inv[red_idx[part_ref] - x_split]
    # 5. phase 2: rows sorted by reduced length (descending, stable) inside windows; rows longer
    #    than `long_rows` become chunks in slices of their own; all-empty slices optionally dropped
    rl = np.diff(roff)
    long_ids = np.flatnonzero(rl > long_rows)
    chunks = [(int(r), o, min(long_rows, int(rl[r]) - o)) for r in long_ids
              for o in range(0, int(rl[r]), long_rows)]
    chunk_slices = [chunks[i:i + 32] for i in range(0, len(chunks), 32)]
    row_slices = []
    for w0 in range(0, nr, window):
        rs = np.arange(w0, min(nr, w0 + window))
        ln = np.where(rl[rs] > long_rows, 0, rl[rs])
        live = np.where(rl[rs] > long_rows, -1, rs)          # long rows leave the window
        o = np.lexsort((np.arange(len(rs)), -ln))
        for s0 in range(0, window, 32):
            sel = o[s0:s0 + 32]
            if len(sel) == 0:
                continue
            if skip_empty and ln[sel[0]] == 0:
                continue
            row_slices.append(dict(rows=[int(live[i]) for i in sel], length=int(ln[sel[0]])))
    return dict(ne=ne, nr=nr, band_entries=band_entries, indices=indices, weights=weights, x_split=x_split,
                n_part=n_part, p1=p1, roff=roff, red_idx=red_idx, red_w=red_w, chunk_slices=chunk_slices,
                row_slices=row_slices, long_ids=long_ids, pre_edges=int(cell_len[pre].sum()),
                skip_empty=skip_empty, rl=rl)


def spmv(L, x, init=0.0, scale=1.0):
    x = np.asarray(x, np.float64)
    B = L["band_entries"]
    pval = np.zeros(L["n_part"] + 1)                         # + the zero the padding slots read
    # phase 1: lane l of a slice walks cell l, G indices per load group, padded with the zero entry
    for sl in L["p1"]:
        bandv = np.append(x[sl["band"] * B:(sl["band"] + 1) * B], 0.0)
        bandv = np.append(bandv, np.zeros(B + 1 - len(bandv)))
        for first, ln, slot in sl["cells"]:
            acc = np.zeros(G)
            for g in range(sl["groups"]):
                for t in range(G):
                    j = g * G + t
                    if j < ln:
                        e = first + j
                        w = 1.0 if L["weights"] is None else L["weights"][e]
                        acc[t] += bandv[L["indices"][e] - sl["band"] * B] * w
                    else:
                        acc[t] += bandv[B]                   # padding: the band's zero entry
            pval[slot] = (acc[0] + acc[1]) + (acc[2] + acc[3])
    gather = lambda i: pval[i - L["x_split"]] if i >= L["x_split"] else x[i]
    y = np.full(L["nr"], np.nan)
    if L["skip_empty"]:
        y[:] = init                                          # the caller clears the output
    # phase 2: chunk slices -> chunk partials; row slices -> epilogue per row
    chunk_partial = {}
    for sl in L["chunk_slices"]:
        for r, o, n in sl:
            a = L["roff"][r] + o
            chunk_partial.setdefault(r, []).append(
                sum(gather(L["red_idx"][a + j]) * L["red_w"][a + j] for j in range(n)))
    for sl in L["row_slices"]:
        for r in sl["rows"]:
            if r < 0:
                continue
            a, n = L["roff"][r], L["rl"][r]
            assert n <= sl["length"]
            y[r] = init + scale * sum(gather(L["red_idx"][a + j]) * L["red_w"][a + j] for j in range(n))
    for r, parts in chunk_partial.items():                  # the tail kernel: fixed order
        y[r] = init + scale * sum(parts)
    return y
