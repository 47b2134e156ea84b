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


def run_task(L, t, x, pval, dtype, minc=3):
    """band_reduce_kernel, one task on one warp: 32 lanes x 8 edges per step (the same
    statements as the CUDA kernel, including the 3-cell-starts-per-lane bound)."""
    bits = L["bits"]
    mask = (1 << bits) - 1
    bidx, bw, ppos = L["bidx"], L["bw"], L["ppos"]
    band0 = t["band"] << bits
    a0 = t["first"] & ~7
    lead = t["first"] - a0
    total = lead + t["count"]
    steps = (total + 255) // 256
    pp = t["rank_first"]
    lane_flags = (8 + minc - 1) // minc
    rank = 0
    carry = dtype(0)

    def store(slot, val):
        assert np.isnan(pval[slot]), "partial written twice"
        pval[slot] = val

    for st in range(steps):
        cnt = [0] * 32; fm = [0] * 32
        vals = [[dtype(0)] * 8 for _ in range(32)]
        for lane in range(32):
            base = st * 256 + lane * 8
            for i in range(8):
                pos = a0 + base + i
                hv = int(bidx[pos]) if pos < bidx.size else 0
                valid = lead <= base + i < total
                if valid:
                    v = dtype(x[band0 + (hv & mask)])
                    if bw is not None:
                        v = dtype(v * bw[pos])
                    vals[lane][i] = v
                    if hv & 0x8000:
                        fm[lane] |= 1 << i
            cnt[lane] = bin(fm[lane]).count("1")
            assert cnt[lane] <= lane_flags, "too many cell starts in a lane"
        popc = lambda m: bin(m).count("1")
        step_flags = sum(cnt)
        fl = sum((cnt[l] > 0) << l for l in range(32))
        run_tail = [dtype(0)] * 32
        caps = [None] * 32
        slots = [None] * 32
        for lane in range(32):
            excl = sum(cnt[:lane])
            rb = rank + excl - 1
            skip = 1 if rb < 0 else 0
            s = [0] * lane_flags
            for k in range(lane_flags):
                if cnt[lane] > k + skip: s[k] = int(ppos[pp + rb + k + skip])
            slots[lane] = (s, skip)
            c = [dtype(0)] * lane_flags
            run = dtype(0)
            for i in range(8):
                if fm[lane] & (1 << i):
                    if i < minc:
                        c[0] = run
                    else:
                        nb = popc(fm[lane] & ((1 << i) - 1))
                        for k in range(lane_flags):
                            if k * minc <= i and nb == k: c[k] = run
                    run = dtype(0)
                run = dtype(run + vals[lane][i])
            caps[lane] = c
            run_tail[lane] = run
        sv = list(run_tail)
        seg0 = []
        for lane in range(32):
            le = fl & ((1 << (lane + 1)) - 1)
            seg0.append(le.bit_length() - 1 if le else 0)
        o = 1
        while o < 32:
            pv = [sv[l - o] if l >= o else dtype(0) for l in range(32)]
            for l in range(32):
                if l - o >= seg0[l]:
                    sv[l] = dtype(sv[l] + pv[l])
            o <<= 1
        for lane in range(32):
            lt = (1 << lane) - 1
            carry_in = carry
            if lane > 0:
                carry_in = sv[lane - 1] if (fl & lt) else dtype(sv[lane - 1] + carry)
            if cnt[lane] > 0:
                c = caps[lane]
                s, skip = slots[lane]
                c = list(c)
                c[0] = dtype(c[0] + carry_in)
                if skip:
                    c = c[1:] + [c[-1]]
                for k in range(lane_flags):
                    if cnt[lane] > k + skip: store(s[k], c[k])
        carry = sv[31] if fl else dtype(sv[31] + carry)
        rank += step_flags
    if rank > 0:
        store(int(ppos[pp + rank - 1]), carry)


def run_task2(L, t, x, pval, dtype, le=16):
    """band_reduce2_kernel, one task on one warp: 32 lanes x `le` edges per step, running sums
    through a scratch column, cells closed by a loop over the lane's cell-start bits."""
    bits = L["bits"]
    mask = (1 << bits) - 1
    bidx, bw, ppos = L["bidx"], L["bw"], L["ppos"]
    band0 = t["band"] << bits
    a0 = t["first"] & ~7
    lead = t["first"] - a0
    total = lead + t["count"]
    se = 32 * le
    steps = (total + se - 1) // se
    pp = t["rank_first"]
    rank = 0
    carry = dtype(0)

    def store(slot, val):
        assert np.isnan(pval[slot]), "partial written twice"
        pval[slot] = val

    for st in range(steps):
        fm = [0] * 32
        scratch = [[dtype(0)] * le for _ in range(32)]
        tails = [dtype(0)] * 32
        for lane in range(32):
            base = st * se + lane * le
            run = dtype(0)
            for i in range(le):
                pos = a0 + base + i
                hv = int(bidx[pos]) if pos < bidx.size else 0
                valid = lead <= base + i < total
                v = dtype(0)
                flag = False
                if valid:
                    v = dtype(x[band0 + (hv & mask)])
                    if bw is not None:
                        v = dtype(v * bw[pos])
                    flag = bool(hv & 0x8000)
                scratch[lane][i] = run
                if flag:
                    fm[lane] |= 1 << i
                    run = dtype(0)
                run = dtype(run + v)
            tails[lane] = run
        cnt = [bin(m).count("1") for m in fm]
        assert max(cnt) <= 7
        step_flags = sum(cnt)
        fl = sum((cnt[l] > 0) << l for l in range(32))
        sv = list(tails)
        seg0 = []
        for lane in range(32):
            le_m = fl & ((1 << (lane + 1)) - 1)
            seg0.append(le_m.bit_length() - 1 if le_m else 0)
        o = 1
        while o < 32:
            pv = [sv[l - o] if l >= o else dtype(0) for l in range(32)]
            for l in range(32):
                if l - o >= seg0[l]:
                    sv[l] = dtype(sv[l] + pv[l])
            o <<= 1
        for lane in range(32):
            lt = (1 << lane) - 1
            carry_in = carry
            if lane > 0:
                carry_in = sv[lane - 1] if (fl & lt) else dtype(sv[lane - 1] + carry)
            r = rank + sum(cnt[:lane]) - 1
            m = fm[lane]
            add = carry_in
            while m:
                i = (m & -m).bit_length() - 1
                m &= m - 1
                val = dtype(scratch[lane][i] + add)
                add = dtype(0)
                if r >= 0:
                    store(int(ppos[pp + r]), val)
                r += 1
        carry = sv[31] if fl else dtype(sv[31] + carry)
        rank += step_flags
    if rank > 0:
        store(int(ppos[pp + rank - 1]), carry)


def spmv_banded(L, x, dtype=np.float64, minc=3, kernel=1, le=16):
    pval = np.full(L["n_part"], np.nan, dtype)
    for t in L["tasks"]:
        if kernel == 2:
            run_task2(L, t, x, pval, dtype, le)
        else:
            run_task(L, t, x, pval, dtype, minc)
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


# ---------------------------------------------------------------------------------------------
# Round-2 layout: a row's direct edges become "global cells", so that
# phase 2 is a streaming segmented sum over partials only.  Model of the layout and of both
# phases; not implemented in CUDA yet.
# ---------------------------------------------------------------------------------------------

def add_global_cells(L, off, idx, w, cellmax=2048, window=2048):
    """On top of build_layout: every edge that is NOT pre-reduced goes, in CSC order, into the
    global stream `gidx` (32-bit source | cell start << 31, weights in `gw`).  A global cell is a
    run of one row's direct edges, cut where the stream position is a multiple of `cellmax`
    (every task window of `window` >= cellmax stream positions then starts at most one cell late).
    A row's partial slots are consecutive: first its banded cells in CSC order, then its global
    cells in stream order; `poff[r] .. poff[r + 1]` delimits them, `ppos` (band-major rank ->
    slot) is renumbered accordingly and `gpos` maps the global cell rank to its slot."""
    off = np.asarray(off, np.int64)
    idx = np.asarray(idx, np.int64)
    ne = idx.size
    nrows = off.size - 1
    banded, f = L["banded"], L["f"]
    row = np.repeat(np.arange(nrows), np.diff(off))
    direct = ~banded
    ge = np.flatnonzero(direct)                    # CSC positions of the stream's edges
    grow = row[ge]
    gstart = np.ones(ge.size, bool)
    gstart[1:] = grow[1:] != grow[:-1]
    gstart |= (np.arange(ge.size) % cellmax) == 0
    gidx = (idx[ge] | (gstart.astype(np.int64) << 31)).astype(np.uint32)
    gw = None if w is None else w[ge]
    head = banded & f                              # banded cell starts, CSC order
    n_bcells = np.bincount(row[head], minlength=nrows)
    n_gcells = np.bincount(grow[gstart], minlength=nrows)
    poff = np.concatenate(([0], np.cumsum(n_bcells + n_gcells)))
    # slot of each banded cell: poff[row] + its rank among the row's banded cells
    hpos = np.flatnonzero(head)
    hrow = row[hpos]
    first_of_row = np.concatenate(([0], np.cumsum(n_bcells)))[hrow]
    slot_b = poff[hrow] + (np.arange(hpos.size) - first_of_row)
    # ppos is indexed by band-major cell rank; L["ppos"] holds the OLD slot = CSC-order cell rank
    ppos = slot_b[L["ppos"]]
    gcell_row = grow[gstart]
    gfirst = np.concatenate(([0], np.cumsum(n_gcells)))[gcell_row]
    gpos = poff[gcell_row] + n_bcells[gcell_row] + (np.arange(gcell_row.size) - gfirst)
    # tasks over the global stream, as for a band: the cells that START inside a window
    starts = np.flatnonzero(gstart)
    gtasks = []
    for lo in range(0, ge.size, window):
        hi = min(lo + window, ge.size)
        a = int(np.searchsorted(starts, lo))
        b = int(np.searchsorted(starts, hi))
        first = int(starts[a]) if a < starts.size else ge.size
        end = int(starts[b]) if b < starts.size else ge.size
        gtasks.append(dict(first=first, rank_first=a, count=max(end - first, 0), band=0))
    G = dict(L)
    G.update(ppos=ppos, gidx=gidx, gw=gw, gpos=gpos, gtasks=gtasks, poff=poff,
             n_slots=int(poff[-1]), n_global=int(ge.size))
    return G


def spmv_global_cells(G, x, dtype=np.float64, le=16):
    """Phase 1: band tasks (band_reduce2 lane algorithm, shared-memory gathers) and global tasks
    (the same lane algorithm with 32-bit indices and global gathers) fill `pval`; phase 2 is
    y[r] = sum(pval[poff[r]:poff[r + 1]])."""
    pval = np.full(G["n_slots"], np.nan, dtype)
    for t in G["tasks"]:
        run_task2(G, t, x, pval, dtype, le)
    # the global stream through the same routine: a "band" of 2^31 entries at offset 0
    S = dict(bits=31, bidx=G["gidx"], bw=G["gw"], ppos=G["gpos"])
    for t in G["gtasks"]:
        _run_task2_wide(S, t, x, pval, dtype, le)
    assert not np.isnan(pval).any(), "a partial was never written"
    poff = G["poff"]
    y = np.zeros(poff.size - 1, dtype)
    nz = poff[:-1] < poff[1:]
    y[nz] = np.add.reduceat(pval, poff[:-1][nz])
    return y


def _run_task2_wide(S, t, x, pval, dtype, le):
    """run_task2 with the cell-start flag in bit 31 of a 32-bit index (global cells)."""
    wide = dict(S)
    flags = (S["bidx"] >> 31).astype(bool)
    # re-express the stream in run_task2's terms: 15-bit-style flag at 0x8000 is too narrow for
    # 31-bit sources, so run the lane algorithm on (source, flag) pairs directly
    src = (S["bidx"] & 0x7FFFFFFF).astype(np.int64)
    a0 = t["first"] & ~7
    lead = t["first"] - a0
    total = lead + t["count"]
    se = 32 * le
    steps = (total + se - 1) // se
    pp = t["rank_first"]
    rank = 0
    carry = dtype(0)
    ppos, bw = wide["ppos"], wide["bw"]

    def store(slot, val):
        assert np.isnan(pval[slot]), "partial written twice"
        pval[slot] = val

    for st in range(steps):
        fm = [0] * 32
        scratch = [[dtype(0)] * le for _ in range(32)]
        tails = [dtype(0)] * 32
        for lane in range(32):
            base = st * se + lane * le
            run = dtype(0)
            for i in range(le):
                pos = a0 + base + i
                valid = lead <= base + i < total
                v = dtype(0)
                flag = False
                if valid:
                    v = dtype(x[src[pos]])
                    if bw is not None:
                        v = dtype(v * bw[pos])
                    flag = bool(flags[pos])
                scratch[lane][i] = run
                if flag:
                    fm[lane] |= 1 << i
                    run = dtype(0)
                run = dtype(run + v)
            tails[lane] = run
        cnt = [bin(m).count("1") for m in fm]
        step_flags = sum(cnt)
        fl = sum((cnt[l] > 0) << l for l in range(32))
        sv = list(tails)
        seg0 = []
        for lane in range(32):
            le_m = fl & ((1 << (lane + 1)) - 1)
            seg0.append(le_m.bit_length() - 1 if le_m else 0)
        o = 1
        while o < 32:
            pv = [sv[l - o] if l >= o else dtype(0) for l in range(32)]
            for l in range(32):
                if l - o >= seg0[l]:
                    sv[l] = dtype(sv[l] + pv[l])
            o <<= 1
        for lane in range(32):
            lt = (1 << lane) - 1
            carry_in = carry
            if lane > 0:
                carry_in = sv[lane - 1] if (fl & lt) else dtype(sv[lane - 1] + carry)
            r = rank + sum(cnt[:lane]) - 1
            m = fm[lane]
            add = carry_in
            while m:
                i = (m & -m).bit_length() - 1
                m &= m - 1
                val = dtype(scratch[lane][i] + add)
                add = dtype(0)
                if r >= 0:
                    store(int(ppos[pp + r]), val)
                r += 1
        carry = sv[31] if fl else dtype(sv[31] + carry)
        rank += step_flags
    if rank > 0:
        store(int(ppos[pp + rank - 1]), carry)
