"""Executable model of the banded pre-reduction (cugraph_forked_b200/csrc/banded.cu + the
phase-1 kernel of banded.cuh), statement by statement in numpy / plain Python, for checking the
algorithm on the CPU: layout build, the lane-level warp algorithm of band_reduce_kernel, and
phase 2 over the reduced CSC.  Not part of the product path and not an oracle of the reference;
tests/test_banded_model.py runs it against a plain CSC SpMV."""
from __future__ import annotations

import numpy as np


def build_layout(off, idx, w, num_minors, bits, minc=3, cellmax=2048, window=2048, hot_min=32768):
    off = np.asarray(off, np.int64)
    idx = np.asarray(idx, np.int64)
    ne = idx.size
    e = np.arange(ne)
    band = idx >> bits
    nb = int((num_minors + (1 << bits) - 1) >> bits)
    # mark_row_starts_kernel / mark_cell_starts_kernel
    rowstart = np.zeros(ne, bool)
    nonempty = off[:-1] < off[1:]
    rowstart[off[:-1][nonempty]] = True
    f = rowstart | (e == 0) | (e % cellmax == 0)
    f[1:] |= band[1:] != band[:-1]
    # classify_kernel: nearest cell start at or before e within minc - 1 edges
    p = np.full(ne, -1)
    for j in range(minc):
        ok = (p < 0) & (e - j >= 0)
        ok[ok] &= f[(e - j)[ok]]
        p[ok] = (e - j)[ok]
    cand = np.ones(ne, bool)
    have = p >= 0
    for j in range(1, minc):
        q = p + j
        bad = have & ((q >= ne) | f[np.minimum(q, ne - 1)])
        cand &= ~bad
    hist = np.bincount(band[cand], minlength=nb)
    hot = hist >= hot_min
    # finalize_masks_kernel
    banded = cand & hot[band]
    keep = ~banded | f
    head = banded & f
    excl = lambda m: np.concatenate(([0], np.cumsum(m)))
    newpos, pid, bpos = excl(keep), excl(head), excl(banded)
    n_red, n_part, n_band = int(newpos[-1]), int(pid[-1]), int(bpos[-1])
    x_split = int(num_minors)
    # fill_reduced_kernel / reduced_offsets_kernel
    ridx = np.zeros(n_red, np.int64)
    ridx[newpos[:-1][keep]] = np.where(banded, x_split + pid[:-1], idx)[keep]
    rw = None
    if w is not None:
        rw = np.zeros(n_red, w.dtype)
        rw[newpos[:-1][keep]] = np.where(banded, 1, w)[keep]
    roff = newpos[off]
    # radix sort by band (stable) + fill_band_major_kernel
    eb = e[banded]
    order = np.argsort(band[eb], kind="stable")
    ej = eb[order]
    bidx = ((idx[ej] & ((1 << bits) - 1)) | (f[ej].astype(np.int64) << 15)).astype(np.uint16)
    bw = w[ej] if w is not None else None
    hf = f[ej]
    rank = excl(hf)
    assert int(rank[-1]) == n_part
    ppos = np.zeros(n_part, np.int64)
    fpos = np.zeros(n_part, np.int64)
    ppos[rank[:-1][hf]] = pid[:-1][ej[hf]]
    fpos[rank[:-1][hf]] = np.arange(n_band)[hf]
    bstart = np.searchsorted(band[ej], np.arange(nb + 1), side="left")
    # tasks (host windows + finalize_tasks_kernel)
    def next_start(pos):
        if pos >= n_band:
            return n_band, n_part
        q = int(rank[pos])
        return (int(fpos[q]) if q < n_part else n_band), q
    tasks = []
    for k in range(nb):
        for lo in range(int(bstart[k]), int(bstart[k + 1]), window):
            hi = min(lo + window, int(bstart[k + 1]))
            a, q0 = next_start(lo)
            b, _ = next_start(hi)
            tasks.append(dict(first=a, rank_first=q0, count=max(b - a, 0), band=k))
    return dict(bits=bits, x_split=x_split, roff=roff, ridx=ridx, rw=rw, bidx=bidx, bw=bw, ppos=ppos,
                tasks=tasks, n_band=n_band, n_part=n_part, n_red=n_red, f=f, banded=banded, cand=cand,
                nb=nb)


def run_task(L, t, x, pval, dtype):
    """band_reduce_kernel, one task on one warp: 32 lanes x 8 edges per step."""
    bits = L["bits"]
    mask = (1 << bits) - 1
    bidx, bw, ppos = L["bidx"], L["bw"], L["ppos"]
    band0 = t["band"] << bits
    a0 = t["first"] & ~7
    lead = t["first"] - a0
    total = lead + t["count"]
    steps = (total + 255) // 256
    rank = t["rank_first"]
    carry = dtype(0)
    for st in range(steps):
        fm = [0] * 32; vm = [0] * 32
        h = [[0] * 8 for _ in range(32)]
        for lane in range(32):
            base = st * 256 + lane * 8
            for i in range(8):
                pos = a0 + base + i
                hv = int(bidx[pos]) if pos < bidx.size else 0
                h[lane][i] = hv
                valid = lead <= base + i < total
                if valid:
                    vm[lane] |= 1 << i
                    if hv & 0x8000:
                        fm[lane] |= 1 << i
        cnt = [bin(m).count("1") for m in fm]
        incl = np.cumsum(cnt)
        excl = incl - cnt
        step_flags = int(incl[31])
        run_tail = [dtype(0)] * 32
        closev = [[dtype(0)] * 8 for _ in range(32)]
        slot = [[-1] * 8 for _ in range(32)]
        for lane in range(32):
            base = st * 256 + lane * 8
            for i in range(8):
                if fm[lane] & (1 << i):
                    r = rank + excl[lane] + bin(fm[lane] & ((1 << i) - 1)).count("1") - 1
                    if r >= t["rank_first"]:
                        slot[lane][i] = int(ppos[r])
            run = dtype(0)
            for i in range(8):
                v = dtype(0)
                if vm[lane] & (1 << i):
                    v = dtype(x[band0 + (h[lane][i] & mask)])
                    if bw is not None:
                        v = dtype(v * bw[a0 + base + i])
                closev[lane][i] = run
                if fm[lane] & (1 << i):
                    run = dtype(0)
                run = dtype(run + v)
            run_tail[lane] = run
        # segmented inclusive scan (Hillis-Steele, exactly the shuffle sequence)
        sv = list(run_tail)
        sf = [c > 0 for c in cnt]
        o = 1
        while o < 32:
            pv = [sv[l - o] if l >= o else dtype(0) for l in range(32)]
            pf = [sf[l - o] if l >= o else False for l in range(32)]
            for l in range(32):
                if l >= o:
                    if not sf[l]:
                        sv[l] = dtype(sv[l] + pv[l])
                    sf[l] = sf[l] or pf[l]
            o <<= 1
        for lane in range(32):
            if lane == 0:
                carry_in = carry
            else:
                carry_in = sv[lane - 1] if sf[lane - 1] else dtype(sv[lane - 1] + carry)
            first = True
            for i in range(8):
                if fm[lane] & (1 << i):
                    val = dtype(closev[lane][i] + carry_in) if first else closev[lane][i]
                    first = False
                    if slot[lane][i] >= 0:
                        assert np.isnan(pval[slot[lane][i]]), "partial written twice"
                        pval[slot[lane][i]] = val
        carry = sv[31] if sf[31] else dtype(sv[31] + carry)
        rank += step_flags
    if rank > t["rank_first"]:
        s = int(ppos[rank - 1])
        assert np.isnan(pval[s]), "partial written twice"
        pval[s] = carry


def spmv_banded(L, x, dtype=np.float64):
    pval = np.full(L["n_part"], np.nan, dtype)
    for t in L["tasks"]:
        run_task(L, t, x, pval, dtype)
    assert not np.isnan(pval).any(), "a partial was never written"
    xcat = np.concatenate([x.astype(dtype), pval])
    prod = xcat[L["ridx"]]
    if L["rw"] is not None:
        prod = prod * L["rw"]
    roff = L["roff"]
    y = np.zeros(roff.size - 1, dtype)
    nz = roff[:-1] < roff[1:]
    if prod.size:
        y[nz] = np.add.reduceat(prod, roff[:-1][nz])
    return y
