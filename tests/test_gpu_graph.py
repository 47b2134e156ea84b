"""GPU: the CUDA graph build (renumber map, degree segments, CSC offsets / indices, weights,
degrees, out-weight sums) is bit-exact against the numpy oracle on the same inputs."""
import numpy as np
import pytest
import torch

from oracle import graph_oracle as go
from tests import util

pytestmark = pytest.mark.gpu


def _build_and_compare(handle, src, dst, w=None, vertices=None, store_transposed=True):
    from cugraph_forked_b200 import engine
    g = engine.Graph(handle, torch.from_numpy(src), torch.from_numpy(dst),
                     None if w is None else torch.from_numpy(w),
                     vertices=None if vertices is None else torch.from_numpy(vertices),
                     store_transposed=store_transposed)
    maj, mnr = (dst, src) if store_transposed else (src, dst)
    labels, seg, nmaj, nmin = go.renumber_sg(maj, mnr, vertices)
    offs, idx, ww = go.compress_sorted(nmaj, nmin, w, 0, len(labels))
    assert g.num_vertices == len(labels) and g.num_edges == len(src)
    assert np.array_equal(g.renumber_map().cpu().numpy(), labels)
    assert g.segment_offsets() == seg
    assert np.array_equal(g.offsets().cpu().numpy().astype(np.int64), offs)
    assert np.array_equal(g.indices().cpu().numpy(), idx.astype(np.int32))
    if w is not None:
        assert np.array_equal(g.weights().cpu().numpy(), ww)
    assert np.array_equal(g.in_degrees().cpu().numpy().astype(np.int64), np.diff(offs))
    if store_transposed:
        ow = go.out_weight_sums(idx, ww, len(labels), np.float64)
        got = g.out_weight_sums(torch.float64).cpu().numpy()
        if w is None:
            assert np.array_equal(got, ow)
        else:
            np.testing.assert_allclose(got, ow, rtol=1e-12)
    return g


@pytest.mark.parametrize("name", ["karate.mtx", "dolphins.mtx", "netscience.mtx"])
@pytest.mark.parametrize("weighted", [False, True])
def test_mtx_graphs(handle, name, weighted):
    src, dst, w, nv, _ = util.load_mtx(name)
    _build_and_compare(handle, src, dst, w if weighted else None)
    # with the full vertex list (isolated vertices kept, as the C++ tests' File_Usecase does)
    _build_and_compare(handle, src, dst, w if weighted else None, np.arange(nv, dtype=np.int32))


@pytest.mark.parametrize("scale", [4, 10, 14, 17])
def test_rmat_graphs(handle, scale):
    src, dst = util.rmat(scale)
    _build_and_compare(handle, src, dst, None, np.arange(1 << scale, dtype=np.int32))
    _build_and_compare(handle, src, dst)  # vertex set = ids present in the edge list


def test_rmat_weighted_and_csr(handle):
    src, dst = util.rmat(12)
    w = np.random.default_rng(1).random(len(src)).astype(np.float32)
    _build_and_compare(handle, src, dst, w)
    _build_and_compare(handle, src, dst, w.astype(np.float64))
    _build_and_compare(handle, src, dst, w, store_transposed=False)


def test_int64_and_sparse_ids(handle):
    src, dst = util.rmat(10, dtype=np.int64)
    _build_and_compare(handle, src, dst)
    # sparse external ids far apart: exercises the sort + unique + binary-search path
    rng = np.random.default_rng(5)
    ext = np.sort(rng.choice(np.int64(1) << 40, size=1 << 10, replace=False)).astype(np.int64)
    _build_and_compare(handle, ext[src], ext[dst])
    _build_and_compare(handle, ext[src], ext[dst], vertices=ext)


def test_generator_bit_exact(handle):
    from cugraph_forked_b200 import engine
    for scale, ne, seed, flip in [(10, 5000, 0, False), (20, 100000, 42, False), (12, 4096, 7, True)]:
        s, d = engine.generate_rmat_edgelist(handle, scale, ne, seed=seed, clip_and_flip=flip,
                                             first_edge=123)
        rs, rd = go.rmat_edgelist(scale, ne, seed=seed, clip_and_flip=flip, first_edge=123)
        assert np.array_equal(s.cpu().numpy(), rs) and np.array_equal(d.cpu().numpy(), rd)
    s, d = engine.generate_rmat_edgelist(handle, 34, 1000, seed=1, dtype=torch.int64)
    rs, rd = go.rmat_edgelist(34, 1000, seed=1, dtype=np.int64)
    assert np.array_equal(s.cpu().numpy(), rs) and np.array_equal(d.cpu().numpy(), rd)


def test_edge_cases(handle):
    from cugraph_forked_b200 import engine
    # empty graph
    e = np.zeros(0, np.int32)
    g = engine.Graph(handle, torch.from_numpy(e), torch.from_numpy(e))
    assert g.num_vertices == 0 and g.num_edges == 0
    # single self loop, multi-edges
    _build_and_compare(handle, np.array([3], np.int32), np.array([3], np.int32))
    _build_and_compare(handle, np.array([1, 1, 1, 2], np.int32), np.array([2, 2, 2, 1], np.int32),
                       np.array([1, 2, 3, 4], np.float32))
    # a star: one row of degree 5000 (high segment), everything else degree 0 or 1
    n = 5001
    src = np.arange(1, n, dtype=np.int32)
    dst = np.zeros(n - 1, np.int32)
    g = _build_and_compare(handle, src, dst)
    assert g.segment_offsets() == [0, 1, 1, 1, n]
    # negative ids are rejected
    with pytest.raises(engine.CugraphError, match="non-negative"):
        engine.Graph(handle, torch.tensor([-1, 2], dtype=torch.int32),
                     torch.tensor([1, 2], dtype=torch.int32))
    # endpoints outside the given vertex list are rejected
    with pytest.raises(engine.CugraphError, match="vertex list"):
        engine.Graph(handle, torch.tensor([0, 9], dtype=torch.int32),
                     torch.tensor([1, 2], dtype=torch.int32),
                     vertices=torch.tensor([0, 1, 2, 5], dtype=torch.int32))


def test_renumber_false(handle):
    from cugraph_forked_b200 import engine
    src, dst = util.rmat(9, edge_factor=4)
    g = engine.Graph(handle, torch.from_numpy(src), torch.from_numpy(dst), renumber=False)
    nv = int(max(src.max(), dst.max())) + 1
    assert g.num_vertices == nv and g.segment_offsets() == []
    assert np.array_equal(g.renumber_map().cpu().numpy(), np.arange(nv, dtype=np.int32))
    offs, idx, _ = go.compress_sorted(dst, src, None, 0, nv)
    assert np.array_equal(g.offsets().cpu().numpy().astype(np.int64), offs)
    assert np.array_equal(g.indices().cpu().numpy(), idx)
