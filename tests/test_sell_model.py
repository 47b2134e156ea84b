"""CPU: the executable model of the round-2 "lane owns a segment" layout
(scripts/analysis/sell_layout_model.py, the rules csrc/sell.cu follows) against a plain CSC SpMV,
plus the layout's invariants: every edge is either pre-reduced or direct exactly once, partial
slots are a permutation in band-major sorted order, slices are sorted by length, long rows are
chunked, skip_empty drops only slices whose rows are all empty."""
import numpy as np
import pytest

from scripts.analysis import sell_layout_model as m
from tests import util


def _csc(scale, seed, weighted):
    src, dst = util.rmat(scale, edge_factor=8, seed=seed)
    nv = 1 << scale
    key = np.lexsort((src, dst))                      # rows = destinations, ascending sources
    offsets = np.zeros(nv + 1, np.int64)
    np.cumsum(np.bincount(dst, minlength=nv), out=offsets[1:])
    w = np.random.default_rng(seed).random(len(src)) if weighted else None
    return offsets, src[key].astype(np.int64), None if w is None else w[key], nv


def _plain(offsets, indices, weights, x):
    rows = np.repeat(np.arange(len(offsets) - 1), np.diff(offsets))
    vals = x[indices] * (1.0 if weights is None else weights)
    return np.bincount(rows, weights=vals, minlength=len(offsets) - 1)


@pytest.mark.parametrize("weighted", [False, True])
@pytest.mark.parametrize("params", [dict(), dict(band_entries=32, cell_max=4, long_rows=3),
                                    dict(min_cell=3, hot_min=40), dict(skip_empty=True)])
def test_model_matches_plain_spmv(weighted, params):
    offsets, indices, weights, nv = _csc(8, 3, weighted)
    L = m.build(offsets, indices, weights, **params)
    x = np.random.default_rng(1).random(nv) + 0.25
    y = m.spmv(L, x, init=0.5, scale=2.0)
    ref = 0.5 + 2.0 * _plain(offsets, indices, weights, x)
    assert not np.isnan(y).any()
    np.testing.assert_allclose(y, ref, rtol=1e-12, atol=0)


def test_layout_invariants():
    offsets, indices, weights, nv = _csc(9, 5, False)
    L = m.build(offsets, indices, None, band_entries=64, cell_max=8, long_rows=6)
    # edges: pre-reduced + direct = all; reduced entries = direct + one per cell
    direct = int((L["red_idx"] < L["x_split"]).sum())
    assert L["pre_edges"] + direct == L["ne"]
    assert len(L["red_idx"]) == direct + L["n_part"]
    # partial slots: every slot referenced exactly once, assigned in slice order
    refs = np.sort(L["red_idx"][L["red_idx"] >= L["x_split"]] - L["x_split"])
    assert np.array_equal(refs, np.arange(L["n_part"]))
    slots = [s for sl in L["p1"] for _, _, s in sl["cells"]]
    assert slots == list(range(L["n_part"]))
    # phase-1 slices: one band each, lengths descending inside a band, cells >= 2 edges, <= cell_max
    prev = {}
    for sl in L["p1"]:
        lens = [l for _, l, _ in sl["cells"]]
        assert lens == sorted(lens, reverse=True) and min(lens) >= 2 and max(lens) <= 8
        assert sl["groups"] == -(-lens[0] // m.G)
        assert prev.get(sl["band"], 10**9) >= lens[0]
        prev[sl["band"]] = lens[-1]
        for f, l, _ in sl["cells"]:
            assert np.all(indices[f:f + l] // 64 == sl["band"])
    # phase 2: every row once (in a row slice, or - if long - in chunks), lengths descending
    seen = np.zeros(nv, int)
    for sl in L["row_slices"]:
        ls = [L["rl"][r] if r >= 0 else 0 for r in sl["rows"]]
        assert ls == sorted(ls, reverse=True) and sl["length"] == ls[0]
        for r in sl["rows"]:
            if r >= 0:
                seen[r] += 1
    for sl in L["chunk_slices"]:
        for r, o, n in sl:
            assert L["rl"][r] > 6 and 0 < n <= 6
    seen[L["long_ids"]] += 1
    assert np.all(seen == 1)


def test_skip_empty_drops_only_empty_slices():
    offsets, indices, weights, nv = _csc(9, 7, False)
    full = m.build(offsets, indices, None, window=64)
    skip = m.build(offsets, indices, None, window=64, skip_empty=True)
    kept = {tuple(s["rows"]) for s in skip["row_slices"]}
    for s in full["row_slices"]:
        if tuple(s["rows"]) not in kept:
            assert s["length"] == 0
    assert len(skip["row_slices"]) < len(full["row_slices"])
