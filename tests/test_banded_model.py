"""CPU check of the banded pre-reduction algorithm (cugraph_forked_b200/csrc/banded.cu[h]) through
its executable model scripts/analysis/banded_model.py: the two-phase result must equal a plain
CSC SpMV, every partial must be written exactly once, and the local cell-size rule must agree
with the true cell sizes.  The CUDA implementation itself is covered by the -m gpu tests."""
import numpy as np
import pytest

from oracle import graph_oracle as go
from scripts.analysis import banded_model as bm


def _csc(scale, ef, seed):
    nv, ne = 1 << scale, ef << scale
    src, dst = go.rmat_edgelist(scale, ne, seed=seed)
    indeg = np.bincount(dst, minlength=nv)
    order = np.argsort(-indeg, kind="stable")
    newid = np.empty(nv, np.int64)
    newid[order] = np.arange(nv)
    s, d = newid[src], newid[dst]
    key = np.sort((d << 32) | s)
    rows, cols = key >> 32, key & 0xFFFFFFFF
    off = np.concatenate(([0], np.cumsum(np.bincount(rows, minlength=nv))))
    return nv, off, cols


@pytest.mark.parametrize("bits,cellmax,window,minc,hot_min,weighted", [
    (6, 64, 64, 3, 16, False),
    (6, 32, 64, 3, 1, True),
    (4, 16, 16, 2, 8, True),
    (6, 64, 64, 2, 16, False),
    (8, 2048, 2048, 3, 64, False),
    (5, 24, 40, 4, 1, False),
])
def test_two_phase_equals_plain_spmv(bits, cellmax, window, minc, hot_min, weighted):
    nv, off, idx = _csc(10, 16, seed=3)
    rng = np.random.default_rng(1)
    x = rng.random(nv)
    w = rng.random(idx.size) if weighted else None
    L = bm.build_layout(off, idx, w, nv, bits, minc=minc, cellmax=cellmax, window=window,
                        hot_min=hot_min)
    assert L["n_band"] > 0 and L["n_part"] > 0
    # the local rule of classify_kernel == true cell size >= minc
    starts = np.flatnonzero(L["f"])
    sizes = np.diff(np.append(starts, idx.size))
    true_cand = np.repeat(sizes >= minc, sizes)
    assert np.array_equal(true_cand, L["cand"])
    y = bm.spmv_banded(L, x, minc=minc)
    prod = x[idx] * (w if weighted else 1.0)
    ref = np.zeros(nv)
    nz = off[:-1] < off[1:]
    ref[nz] = np.add.reduceat(prod, off[:-1][nz])
    assert np.allclose(y, ref, rtol=1e-12, atol=0)
    if minc >= 3:
        for le in (8, 16):
            y2 = bm.spmv_banded(L, x, minc=minc, kernel=2, le=le)
            assert np.allclose(y2, ref, rtol=1e-12, atol=0)
    # every task starts at a cell start and tasks tile the banded edges
    pos = 0
    for t in L["tasks"]:
        if t["count"]:
            assert t["first"] == pos and (L["bidx"][pos] & 0x8000)
            pos += t["count"]
    assert pos == L["n_band"]


def test_float32_error_small():
    nv, off, idx = _csc(10, 16, seed=5)
    rng = np.random.default_rng(2)
    x = rng.random(nv).astype(np.float32)
    L = bm.build_layout(off, idx, None, nv, 6, cellmax=64, window=64, hot_min=16)
    y = bm.spmv_banded(L, x, dtype=np.float32)
    ref = np.zeros(nv)
    nz = off[:-1] < off[1:]
    ref[nz] = np.add.reduceat(x[idx].astype(np.float64), off[:-1][nz])
    assert np.max(np.abs(y - ref) / np.maximum(ref, 1e-30)) < 2e-6


@pytest.mark.parametrize("bits,cellmax,window,hot_min,weighted", [
    (6, 64, 64, 16, False),
    (6, 32, 64, 1, True),
    (8, 2048, 2048, 64, False),
    (5, 24, 40, 1, True),
])
def test_global_cells_layout_equals_plain_spmv(bits, cellmax, window, hot_min, weighted):
    """The round-2 layout: direct edges as global cells, phase 2 a
    streaming segmented sum over partials only."""
    nv, off, idx = _csc(10, 16, seed=7)
    rng = np.random.default_rng(3)
    x = rng.random(nv)
    w = rng.random(idx.size) if weighted else None
    L = bm.build_layout(off, idx, w, nv, bits, cellmax=cellmax, window=window, hot_min=hot_min)
    G = bm.add_global_cells(L, off, idx, w, cellmax=cellmax, window=window)
    assert G["n_global"] + G["n_band"] == idx.size and G["n_global"] > 0
    # a row's slots are consecutive and every slot belongs to exactly one cell
    slots = np.concatenate([G["ppos"], G["gpos"]])
    assert np.array_equal(np.sort(slots), np.arange(G["n_slots"]))
    # every global task starts at a cell start and the tasks tile the stream
    pos = 0
    for t in G["gtasks"]:
        if t["count"]:
            assert t["first"] == pos and (int(G["gidx"][pos]) >> 31)
            pos += t["count"]
    assert pos == G["n_global"]
    for le in (8, 16):
        y = bm.spmv_global_cells(G, x, le=le)
        prod = x[idx] * (w if weighted else 1.0)
        ref = np.zeros(nv)
        nz = off[:-1] < off[1:]
        ref[nz] = np.add.reduceat(prod, off[:-1][nz])
        assert np.allclose(y, ref, rtol=1e-12, atol=0)
