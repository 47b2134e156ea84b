"""GPU parity of HITS (cugraph_forked_b200/csrc/hits.cu; SURVEY.md 8f rank 1, the first row after
the hot path) through the raw C ABI: the golden vectors of cpp/tests/c_api/hits_test.c and the
CPU oracle (oracle/centrality_oracle.cpp hits_cpu, itself pinned bit-identical to the reference's
hits_reference) at an identical iteration count.  Tolerances as for the other drivers: 1e-5
(fp32) / 1e-12 (fp64) per vertex relative, identical top-100.  The file sorts last on purpose: it
is the newest code of the round."""
import numpy as np
import pytest
import torch

from oracle import cpu_oracle as co
from tests import golden_vectors as gv
from tests import util
from tests.test_gpu_centrality import TDT, _graph

pytestmark = pytest.mark.gpu


def _ext(t, labels, nv):
    return util.to_external(t.cpu().numpy(), labels, nv)


def test_hits_golden_vectors(handle, oracle_built):
    from cugraph_forked_b200 import engine
    # the C test creates the graph WITH weights; HITS ignores them (hits_impl.cuh:114-140)
    g, labels = _graph(handle, gv.G6_SRC, gv.G6_DST, gv.G6_WGT)
    p = gv.G6_HITS
    off, idx = co.csr_from_coo(gv.G6_SRC, gv.G6_DST, 6)
    _, _, _, it_ref = co.hits(off, idx, dtype=np.float32, max_iterations=p["max_iterations"],
                              tolerance=p["epsilon"], normalize=False)
    hubs, auth, st = engine.hits(g, epsilon=p["epsilon"], max_iterations=p["max_iterations"],
                                 normalize=False)
    assert st.converged and st.iterations == it_ref
    assert gv.nearly_equal(_ext(hubs, labels, 6), p["hubs"])
    assert gv.nearly_equal(_ext(auth, labels, 6), p["authorities"])
    # initial hub guess (hits_test.c:221-253), given in internal order
    p = gv.G6_HITS_INITIAL
    start = np.zeros(6, np.float32)
    start[p["initial_vertices"]] = p["initial_hubs"]
    hubs, auth, _ = engine.hits(g, epsilon=p["epsilon"], max_iterations=p["max_iterations"],
                                normalize=False, initial_hubs_guess=torch.from_numpy(start[labels]))
    assert gv.nearly_equal(_ext(hubs, labels, 6), p["hubs"])
    assert gv.nearly_equal(_ext(auth, labels, 6), p["authorities"])
    # the 54-vertex graph (hits_test.c:255-300), unweighted
    g, labels = _graph(handle, gv.H54_SRC, gv.H54_DST, None, 54)
    p = gv.H54_HITS
    for dtype in (torch.float32, torch.float64):
        hubs, auth, _ = engine.hits(g, epsilon=p["epsilon"], max_iterations=p["max_iterations"],
                                    normalize=False, dtype=dtype)
        assert gv.nearly_equal(_ext(hubs, labels, 54), p["hubs"])
        assert gv.nearly_equal(_ext(auth, labels, 54), p["authorities"])


@pytest.mark.parametrize("case", ["karate", "rmat12", "rmat16"])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_hits_matches_oracle(handle, oracle_built, case, dtype):
    from cugraph_forked_b200 import engine
    if case == "karate":
        src, dst, _, nv, _ = util.load_mtx("karate.mtx", dtype)
    else:
        scale = int(case[4:])
        src, dst = util.rmat(scale, seed=6)
        nv = 1 << scale
    g, labels = _graph(handle, src, dst, None, nv)
    off, idx = co.csr_from_coo(src, dst, nv)
    k = 12
    rh, ra, _, _ = co.hits(off, idx, dtype=np.float64, normalize=True, fixed_iterations=k)
    hubs, auth, st = engine.hits(g, dtype=TDT[dtype], normalize=True, fixed_iterations=k)
    tol = 1e-5 if dtype == np.float32 else 1e-12
    for got, ref in ((_ext(hubs, labels, nv), rh), (_ext(auth, labels, nv), ra)):
        assert np.all(got[ref == 0] == 0)
        # fp32 leaves the normal range long before fp64 does: skip what would be denormal
        nz = ref > (1e-25 if dtype == np.float32 else 0.0)
        assert util.rel_err(got[nz], ref[nz]).max() <= tol
        assert np.array_equal(util.top_k(got, 100), util.top_k(ref, 100)) or \
            util.rel_err(ref[util.top_k(got, 100)], ref[util.top_k(ref, 100)]).max() <= tol
        assert abs(got.sum() - 1.0) < 1e-5
    # converged run: same iteration count as the oracle of the same precision (fp64), and
    # run-to-run determinism
    if dtype == np.float64:
        _, _, err, it = co.hits(off, idx, dtype=np.float64, tolerance=1e-8, max_iterations=1000)
        h1, a1, st = engine.hits(g, dtype=torch.float64, epsilon=1e-8, max_iterations=1000)
        # the difference of two nearly equal vectors carries ~1 ulp of noise per vertex
        assert st.converged and st.iterations == it
        assert abs(st.last_diff - err) <= 1e-6 * err + 4e-16 * nv
        h2, a2, _ = engine.hits(g, dtype=torch.float64, epsilon=1e-8, max_iterations=1000)
        assert torch.equal(h1, h2) and torch.equal(a1, a2)


def test_hits_errors(handle):
    from cugraph_forked_b200 import engine
    src, dst = util.rmat(8)
    g, _ = _graph(handle, src, dst, None, 1 << 8)
    with pytest.raises(engine.CugraphError, match="epsilon should be non-negative"):
        engine.hits(g, epsilon=-1.0)
    g2 = engine.Graph(handle, torch.from_numpy(src), torch.from_numpy(dst), store_transposed=False)
    with pytest.raises(engine.CugraphError, match="stored transposed"):
        engine.hits(g2)


@pytest.mark.parametrize("store_transposed", [True, False])
@pytest.mark.parametrize("renumber", [True, False])
def test_hits_c_api_golden(store_transposed, renumber):
    """cugraph_hits through the pylibcugraph mirror: hits_test.c test_hits (:155),
    test_hits_with_transpose (:187, store_transposed = FALSE forces the in-place transpose),
    test_hits_with_initial (:221) and test_hits_bigger (:255)."""
    from cugraph_forked_b200 import pylibcugraph as plc
    from tests.test_gpu_c_api import _by_vertex, _sg, _t
    rh = plc.ResourceHandle()
    g = _sg(plc, rh, gv.G6_SRC, gv.G6_DST, gv.G6_WGT, store_transposed, renumber)
    p = gv.G6_HITS
    vertices, hubs, auth = plc.hits(rh, g, p["epsilon"], p["max_iterations"], None, None, False, False)
    assert gv.nearly_equal(_by_vertex(vertices, hubs, 6), p["hubs"])
    assert gv.nearly_equal(_by_vertex(vertices, auth, 6), p["authorities"])
    p = gv.G6_HITS_INITIAL
    vertices, hubs, auth = plc.hits(rh, g, p["epsilon"], p["max_iterations"],
                                    _t(p["initial_vertices"]),
                                    _t(p["initial_hubs"].astype(np.float32)), False, False)
    assert gv.nearly_equal(_by_vertex(vertices, hubs, 6), p["hubs"])
    assert gv.nearly_equal(_by_vertex(vertices, auth, 6), p["authorities"])
    # normalize = TRUE: scores sum to 1 and keep their proportions
    vertices, hubs_n, auth_n = plc.hits(rh, g, gv.G6_HITS["epsilon"], 20, None, None, True, False)
    h = _by_vertex(vertices, hubs_n, 6)
    assert abs(h.sum() - 1.0) < 1e-6 and gv.nearly_equal(h / h.max(), gv.G6_HITS["hubs"])
    # the vertex list of a non-renumbered graph spans 0..max id, so H54 needs renumbering off too
    g = _sg(plc, rh, gv.H54_SRC, gv.H54_DST, np.ones(48), store_transposed, False)
    p = gv.H54_HITS
    vertices, hubs, auth = plc.hits(rh, g, p["epsilon"], p["max_iterations"], None, None, False, False)
    assert gv.nearly_equal(_by_vertex(vertices, hubs, 54), p["hubs"])
    assert gv.nearly_equal(_by_vertex(vertices, auth, 54), p["authorities"])
