"""GPU: the cugraph_c C ABI (through the pylibcugraph-shaped mirror) against the golden vectors
of the reference's own C-API tests and against the oracle.

Follows cpp/tests/c_api/pagerank_test.c:35-318, katz_test.c:34-129,
eigenvector_centrality_test.c:34-110 and create_graph_test.c (error paths)."""
import numpy as np
import pytest
import torch

from tests import golden_vectors as gv
from tests import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def plc():
    from cugraph_forked_b200 import pylibcugraph
    return pylibcugraph


@pytest.fixture(scope="module")
def rh(plc):
    return plc.ResourceHandle()


def _t(a):
    return torch.as_tensor(np.ascontiguousarray(a)).cuda()


def _by_vertex(vertices, values, nv):
    v = vertices.cpu().numpy()
    out = np.full(nv, np.nan)
    out[v] = values.cpu().numpy()
    assert sorted(v.tolist()) == list(range(nv))
    return out


def _sg(plc, rh, src, dst, wgt, store_transposed, renumber=False, dtype=np.float32):
    props = plc.GraphProperties(is_symmetric=False, is_multigraph=False)
    w = None if wgt is None else _t(np.asarray(wgt, dtype))
    return plc.SGGraph(rh, props, _t(src), _t(dst), w, store_transposed=store_transposed,
                       renumber=renumber, do_expensive_check=False)


@pytest.mark.parametrize("store_transposed", [True, False])
@pytest.mark.parametrize("renumber", [True, False])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_pagerank_golden(plc, rh, store_transposed, renumber, dtype):
    g = _sg(plc, rh, gv.G6_SRC, gv.G6_DST, gv.G6_WGT, store_transposed, renumber, dtype)
    p = gv.G6_PAGERANK
    vertices, values = plc.pagerank(rh, g, None, None, None, None, p["alpha"], p["epsilon"],
                                    p["max_iterations"], False)
    assert values.dtype == (torch.float32 if dtype == np.float32 else torch.float64)
    assert gv.nearly_equal(_by_vertex(vertices, values, 6), p["result"])


def test_pagerank_4_golden(plc, rh):
    g = _sg(plc, rh, gv.P4_SRC, gv.P4_DST, gv.P4_WGT, True)
    p = gv.P4_PAGERANK
    vertices, values = plc.pagerank(rh, g, None, None, None, None, p["alpha"], p["epsilon"],
                                    p["max_iterations"], False)
    assert gv.nearly_equal(_by_vertex(vertices, values, 4), p["result"])


@pytest.mark.parametrize("store_transposed", [True, False])
def test_personalized_pagerank_golden(plc, rh, store_transposed):
    g = _sg(plc, rh, gv.P4_SRC, gv.P4_DST, gv.P4_WGT, store_transposed, renumber=True)
    p = gv.P4_PERSONALIZED
    vertices, values = plc.personalized_pagerank(
        rh, g, None, None, None, None, _t(p["vertices"]), _t(p["values"]), p["alpha"],
        p["epsilon"], p["max_iterations"], False)
    assert gv.nearly_equal(_by_vertex(vertices, values, 4), p["result"])


@pytest.mark.parametrize("store_transposed", [True, False])
def test_katz_golden(plc, rh, store_transposed):
    g = _sg(plc, rh, gv.G6_SRC, gv.G6_DST, gv.G6_WGT, store_transposed, renumber=True)
    p = gv.G6_KATZ
    vertices, values = plc.katz_centrality(rh, g, None, p["alpha"], p["beta"], p["epsilon"],
                                           p["max_iterations"], False)
    assert gv.nearly_equal(_by_vertex(vertices, values, 6), p["result"])


def test_katz_betas_are_ignored_like_the_reference(plc, rh):
    """c_api/katz.cpp:158 forwards nullptr for betas whatever the caller passes."""
    g = _sg(plc, rh, gv.G6_SRC, gv.G6_DST, gv.G6_WGT, True, renumber=True)
    p = gv.G6_KATZ
    betas = _t(np.arange(6, dtype=np.float32) + 5)
    vertices, values = plc.katz_centrality(rh, g, betas, p["alpha"], p["beta"], p["epsilon"],
                                           p["max_iterations"], False)
    assert gv.nearly_equal(_by_vertex(vertices, values, 6), p["result"])


@pytest.mark.parametrize("store_transposed", [True, False])
def test_eigenvector_golden(plc, rh, store_transposed):
    g = _sg(plc, rh, gv.E6_SRC, gv.E6_DST, gv.E6_WGT, store_transposed, renumber=True)
    p = gv.E6_EIGENVECTOR
    vertices, values = plc.eigenvector_centrality(rh, g, p["epsilon"], p["max_iterations"], False)
    assert gv.nearly_equal(_by_vertex(vertices, values, 6), p["result"])


def test_pagerank_precomputed_and_initial_guess(plc, rh, oracle_built):
    """Out-weight sums and a warm start given as (vertex, value) pairs with EXTERNAL ids in a
    shuffled order, as collect_local_vertex_values_from_ext_vertex_value_pairs expects."""
    rng = np.random.default_rng(5)
    nv, ne = 500, 6000
    src = rng.integers(0, nv, ne).astype(np.int32)
    dst = rng.integers(0, nv, ne).astype(np.int32)
    wgt = rng.random(ne) + 0.1
    g = _sg(plc, rh, src, dst, wgt, True, renumber=True, dtype=np.float64)
    ow = np.zeros(nv)
    np.add.at(ow, src, wgt)
    perm = rng.permutation(nv).astype(np.int32)
    ref, iters = util.oracle_pagerank(src, dst, wgt, nv, dtype=np.float64, alpha=0.85,
                                      epsilon=1e-10, max_iterations=500)
    guess = np.full(nv, 1.0 / nv)
    vertices, values = plc.pagerank(rh, g, _t(perm), _t(ow[perm]), _t(perm), _t(guess[perm]),
                                    0.85, 1e-10, 500, False)
    got = _by_vertex(vertices, values, nv)
    assert util.rel_err(got, ref).max() <= 1e-9
    # warm start from the converged answer: still the same fixed point
    vertices, values = plc.pagerank(rh, g, None, None, _t(perm), _t(ref[perm]), 0.85, 1e-10, 500,
                                    False)
    assert util.rel_err(_by_vertex(vertices, values, nv), ref).max() <= 1e-8


def test_int64_vertices(plc, rh):
    g = _sg(plc, rh, gv.G6_SRC.astype(np.int64), gv.G6_DST.astype(np.int64), gv.G6_WGT, True,
            renumber=True)
    p = gv.G6_PAGERANK
    vertices, values = plc.pagerank(rh, g, None, None, None, None, p["alpha"], p["epsilon"],
                                    p["max_iterations"], False)
    assert vertices.dtype == torch.int64
    assert gv.nearly_equal(_by_vertex(vertices, values, 6), p["result"])


def test_error_paths(plc, rh):
    props = plc.GraphProperties()
    src, dst = _t(gv.G6_SRC), _t(gv.G6_DST)
    with pytest.raises(ValueError, match="CUGRAPH_INVALID_INPUT.*src size != dst size"):
        plc.SGGraph(rh, props, src, dst[:-1].contiguous(), None, store_transposed=True)
    with pytest.raises(ValueError, match="src type != dst type"):
        plc.SGGraph(rh, props, src, dst.to(torch.int64), None, store_transposed=True)
    with pytest.raises(ValueError, match="src size != weights size"):
        plc.SGGraph(rh, props, src, dst, _t(gv.G6_WGT[:3]), store_transposed=True)
    with pytest.raises(TypeError, match="__cuda_array_interface__"):
        plc.SGGraph(rh, props, gv.G6_SRC, dst, None, store_transposed=True)
    g = _sg(plc, rh, gv.G6_SRC, gv.G6_DST, gv.G6_WGT, True, renumber=True)
    # value type must match the graph's weight type (c_api/pagerank.cpp:256-303)
    with pytest.raises(ValueError, match="must match"):
        plc.pagerank(rh, g, None, None, _t(np.arange(6, dtype=np.int32)),
                     _t(np.full(6, 1 / 6, np.float64)), 0.85, 1e-6, 100, False)
    # personalization vertex that is not in the graph
    with pytest.raises(ValueError):
        plc.personalized_pagerank(rh, g, None, None, None, None,
                                  _t(np.array([77], np.int32)), _t(np.array([1.0], np.float32)),
                                  0.85, 1e-6, 100, False)


def test_resource_handle_rank(plc, rh):
    assert rh.get_rank() == 0
