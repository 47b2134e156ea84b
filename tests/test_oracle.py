"""CPU: pins the oracle (oracle/centrality_oracle.cpp) against the reference's own golden
vectors and against the reference's own host functions compiled from /root/reference
(oracle/_ref, built by oracle/build_ref.sh when the reference tree is present)."""
import numpy as np
import pytest

from oracle import cpu_oracle
from tests import golden_vectors as gv
from tests import util


@pytest.fixture(scope="module", autouse=True)
def _build(oracle_built):
    return oracle_built


IMPLS = ["port"] + (["ref"] if cpu_oracle.ref_available() else [])


@pytest.mark.parametrize("impl", IMPLS)
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_pagerank_c_api_golden_weighted(impl, dtype):
    g = gv.G6_PAGERANK
    pr, it = util.oracle_pagerank(gv.G6_SRC, gv.G6_DST, gv.G6_WGT, 6, dtype=dtype, alpha=g["alpha"],
                                  epsilon=g["epsilon"], max_iterations=g["max_iterations"], impl=impl)
    assert gv.nearly_equal(pr, g["result"])


@pytest.mark.parametrize("impl", IMPLS)
def test_pagerank_c_api_golden_path(impl):
    g = gv.P4_PAGERANK
    pr, _ = util.oracle_pagerank(gv.P4_SRC, gv.P4_DST, gv.P4_WGT, 4, alpha=g["alpha"],
                                 epsilon=g["epsilon"], max_iterations=g["max_iterations"], impl=impl)
    assert gv.nearly_equal(pr, g["result"])
    np.testing.assert_allclose(pr, g["result"], rtol=2e-6)


@pytest.mark.parametrize("impl", IMPLS)
def test_personalized_pagerank_c_api_golden(impl):
    g = gv.P4_PERSONALIZED
    pr, _ = util.oracle_pagerank(gv.P4_SRC, gv.P4_DST, gv.P4_WGT, 4, alpha=g["alpha"],
                                 epsilon=g["epsilon"], max_iterations=g["max_iterations"],
                                 personalization=(g["vertices"], g["values"]), impl=impl)
    assert gv.nearly_equal(pr, g["result"])


@pytest.mark.parametrize("impl", IMPLS)
def test_katz_c_api_golden(impl):
    g = gv.G6_KATZ
    kc, _ = util.oracle_katz(gv.G6_SRC, gv.G6_DST, gv.G6_WGT, 6, alpha=g["alpha"], beta=g["beta"],
                             epsilon=g["epsilon"], max_iterations=g["max_iterations"],
                             normalize=True, impl=impl)
    assert gv.nearly_equal(kc, g["result"])


@pytest.mark.parametrize("impl", IMPLS)
def test_eigenvector_c_api_golden(impl):
    g = gv.E6_EIGENVECTOR
    # the reference oracle accumulates into tmp[src] from old[dst]: feed it the transposed list
    ev, _ = cpu_oracle.eigenvector(gv.E6_DST, gv.E6_SRC, gv.E6_WGT, num_vertices=6,
                                   epsilon=g["epsilon"], max_iterations=g["max_iterations"],
                                   impl=impl)
    assert gv.nearly_equal(ev, g["result"])


@pytest.mark.skipif(not cpu_oracle.ref_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("name", ["karate.mtx", "dolphins.mtx", "netscience.mtx"])
@pytest.mark.parametrize("weighted", [False, True])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_port_is_bit_identical_to_reference_host_functions(name, weighted, dtype):
    src, dst, w, nv, _ = util.load_mtx(name, dtype)
    w = w if weighted else None
    a, ia = util.oracle_pagerank(src, dst, w, nv, dtype=dtype, impl="port")
    b, _ = util.oracle_pagerank(src, dst, w, nv, dtype=dtype, impl="ref")
    assert np.array_equal(a, b)
    rng = np.random.default_rng(7)
    pv = rng.permutation(nv)[: nv // 2].astype(np.int32)
    pw = rng.random(nv // 2).astype(dtype)
    a, _ = util.oracle_pagerank(src, dst, w, nv, dtype=dtype, personalization=(pv, pw), impl="port")
    b, _ = util.oracle_pagerank(src, dst, w, nv, dtype=dtype, personalization=(pv, pw), impl="ref")
    assert np.array_equal(a, b)
    a, _ = util.oracle_katz(src, dst, w, nv, dtype=dtype, alpha=0.01, impl="port")
    b, _ = util.oracle_katz(src, dst, w, nv, dtype=dtype, alpha=0.01, impl="ref")
    assert np.array_equal(a, b)
    a, _ = cpu_oracle.eigenvector(dst, src, w, num_vertices=nv, dtype=dtype, impl="port")
    b, _ = cpu_oracle.eigenvector(dst, src, w, num_vertices=nv, dtype=dtype, impl="ref")
    assert np.array_equal(a, b)


def test_karate_known_values():
    """BASELINE.md section 3 / NetworkX: pr[0] = 0.0970, pr[33] = 0.1009, sum = 1."""
    src, dst, _, nv, sym = util.load_mtx("karate.mtx")
    assert nv == 34 and len(src) == 156 and sym
    pr, it = util.oracle_pagerank(src, dst, None, nv)
    assert abs(pr.sum() - 1.0) < 1e-5
    assert abs(pr[0] - 0.09699753) < 1e-6 and abs(pr[33] - 0.10091895) < 1e-6


def test_fixed_iterations_and_nonconvergence():
    src, dst, _, nv, _ = util.load_mtx("karate.mtx")
    a, ia = util.oracle_pagerank(src, dst, None, nv, dtype=np.float64)
    b, ib = util.oracle_pagerank(src, dst, None, nv, dtype=np.float64, fixed_iterations=ia)
    assert ia == ib and np.array_equal(a, b)
    with pytest.raises(RuntimeError, match="failed to converge"):
        util.oracle_pagerank(src, dst, None, nv, max_iterations=3)


def test_mtx_loader_matches_reference_expansion():
    """symmetric files: every off-diagonal entry is followed by its mirror
    (matrix_market_file_utilities.cu:166-215)."""
    src, dst, w, nv, sym = util.load_mtx("netscience.mtx")
    assert sym and nv == 1589 and len(src) == 5484
    assert np.array_equal(src[0::2], dst[1::2]) and np.array_equal(dst[0::2], src[1::2])
    assert np.array_equal(w[0::2], w[1::2])


# ----------------------------------------------------------------------------------------------
# HITS (the next row): the restatement is pinned to the reference's own
# hits_reference (bit-identical when sum-normalised) and to the golden vectors of
# cpp/tests/c_api/hits_test.c (max-normalised, relative 1e-3)
# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_hits_port_is_bit_identical_to_reference(oracle_built, dtype):
    from oracle import cpu_oracle as co
    if not co.ref_available():
        pytest.skip("oracle/_ref not built")
    cases = []
    for name in ("karate.mtx", "dolphins.mtx", "netscience.mtx"):
        src, dst, _, nv, _ = util.load_mtx(name, dtype)
        cases.append((src, dst, nv))
    src, dst = util.rmat(10, seed=4)
    cases.append((src, dst, 1 << 10))
    rng = np.random.default_rng(5)
    for src, dst, nv in cases:
        off, idx = co.csr_from_coo(src, dst, nv)
        for start in (None, rng.random(nv).astype(dtype)):
            a = co.hits(off, idx, dtype=dtype, max_iterations=200, tolerance=1e-6,
                        starting_hubs=start, impl="ref")
            b = co.hits(off, idx, dtype=dtype, max_iterations=200, tolerance=1e-6,
                        starting_hubs=start, impl="port")
            assert a[3] == b[3] and a[2] == b[2]
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_hits_golden_vectors(oracle_built):
    from oracle import cpu_oracle as co
    off, idx = co.csr_from_coo(gv.G6_SRC, gv.G6_DST, 6)
    p = gv.G6_HITS
    hubs, auth, _, it = co.hits(off, idx, dtype=np.float64, max_iterations=p["max_iterations"],
                                tolerance=p["epsilon"], normalize=False)
    assert it < p["max_iterations"]
    assert gv.nearly_equal(hubs, p["hubs"]) and gv.nearly_equal(auth, p["authorities"])
    p = gv.G6_HITS_INITIAL
    start = np.zeros(6)
    start[p["initial_vertices"]] = p["initial_hubs"]
    hubs, auth, _, _ = co.hits(off, idx, dtype=np.float64, max_iterations=p["max_iterations"],
                               tolerance=p["epsilon"], starting_hubs=start, normalize=False)
    assert gv.nearly_equal(hubs, p["hubs"]) and gv.nearly_equal(auth, p["authorities"])
    off, idx = co.csr_from_coo(gv.H54_SRC, gv.H54_DST, 54)
    p = gv.H54_HITS
    for dtype in (np.float32, np.float64):
        hubs, auth, _, _ = co.hits(off, idx, dtype=dtype, max_iterations=p["max_iterations"],
                                   tolerance=p["epsilon"], normalize=False)
        assert gv.nearly_equal(hubs, p["hubs"]) and gv.nearly_equal(auth, p["authorities"])


@pytest.mark.parametrize("name", sorted(gv.KATZ_TOP10))
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_katz_top10_golden_lists(oracle_built, name, dtype):
    """The golden result files of the reference's legacy Katz test (datasets/ref/katz/*.csv,
    legacy/katz_centrality_test.cu:114-183)."""
    src, dst, _, nv, _ = util.load_mtx(name)
    alpha = 1.0 / (np.bincount(src, minlength=nv).max() + 1)
    for impl in ("port", "ref") if oracle_built.ref_available() else ("port",):
        offsets, indices, _ = oracle_built.csc_from_coo(src, dst, None, nv)
        kc, _ = oracle_built.katz(offsets, indices, None, dtype=dtype, alpha=alpha, beta=1.0,
                                  epsilon=1e-6, max_iterations=100, normalize=True, impl=impl)
        assert util.top_k(kc, 10).tolist() == gv.KATZ_TOP10[name]


@pytest.mark.parametrize("name", sorted(gv.PLC_PAGERANK_RESULTS))
def test_pylibcugraph_pagerank_golden_vectors(oracle_built, name):
    """python/pylibcugraph/pylibcugraph/tests/test_pagerank.py:22-132."""
    p = gv.PLC_PAGERANK
    if name == "Simple_2":
        src, dst, w, nv = gv.G6_SRC, gv.G6_DST, gv.G6_WGT, 6
    else:
        src, dst, w, nv, _ = util.load_mtx(name)
    for impl in ("port", "ref") if oracle_built.ref_available() else ("port",):
        offsets, indices, weights = oracle_built.csc_from_coo(src, dst, w, nv)
        pr, _ = oracle_built.pagerank(offsets, indices, weights, dtype=np.float32, alpha=p["alpha"],
                                      epsilon=p["epsilon"], max_iterations=p["max_iterations"],
                                      impl=impl)
        want = gv.PLC_PAGERANK_RESULTS[name]
        assert np.all(np.abs(pr - want) <= p["rel"] * np.abs(want)), impl


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_reference_pinned_iterations_equals_port(dtype):
    """cpu_oracle.pagerank(impl="ref", fixed_iterations=k): the reference's pagerank_reference run
    with epsilon 0 and max_iterations k does exactly k iterations; bit-identical to the port."""
    if not cpu_oracle.ref_available():
        pytest.skip("oracle/_ref not built")
    src, dst = util.rmat(12, seed=5)
    o, i, _ = cpu_oracle.csc_from_coo(src, dst, None, 1 << 12)
    for k in (1, 7):
        a, _ = cpu_oracle.pagerank(o, i, None, dtype=dtype, impl="ref", fixed_iterations=k)
        b, _ = cpu_oracle.pagerank(o, i, None, dtype=dtype, impl="port", fixed_iterations=k)
        assert np.array_equal(a, b)
