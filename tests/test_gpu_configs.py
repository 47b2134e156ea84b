"""GPU parity on the BASELINE.json configurations themselves: the
default (banded) PageRank path against the reference's own host `pagerank_reference`
(oracle/_ref, compiled from cpp/tests/link_analysis/pagerank_test.cpp:42-130) and against the
fp64 port at the pinned iteration count, on R-MAT scale 20 (configs[1], f32 and f64) and scale 24
(configs[2], f32, a fixed small iteration count).

Asserted: per-vertex relative error <= 1e-5 (f32) / 1e-12 (f64) against the fp64 oracle at the
GPU's iteration count, identical top-100; the reference's own bound (pagerank_test.cpp:355-362)
against `_ref` of the same precision.  Reported (printed, not asserted): max relative error and
top-100 overlap against the fp32 `_ref` - its sequential fp32 row sums carry O(deg * eps) error
on hub rows themselves.

The edge list comes from the library's generator (bit-identical to the oracle's numpy twin,
tests/test_gpu_graph.py) and is copied to the host for the oracle: same inputs on both sides.
"""
import numpy as np
import pytest
import torch

from tests import util

pytestmark = pytest.mark.gpu

TOL = {np.float32: 1e-5, np.float64: 1e-12}
TDT = {np.float32: torch.float32, np.float64: torch.float64}


def _same_top(a, b, k=100, tol=1e-5):
    """identical top-k ranking, up to swaps between vertices whose oracle scores tie within tol"""
    ta, tb = util.top_k(a, k), util.top_k(b, k)
    if np.array_equal(ta, tb):
        return True
    if set(ta[:k - 1]) - set(util.top_k(b, k + 5)):
        return False
    bad = ta != tb
    return bool(np.all(util.rel_err(b[ta[bad]], b[tb[bad]]) <= tol))


def _config(handle, scale):
    from cugraph_forked_b200 import engine
    from oracle import cpu_oracle
    nv, ne = 1 << scale, 16 << scale
    s, d = engine.generate_rmat_edgelist(handle, scale, ne, seed=0)
    g = engine.Graph(handle, s, d, vertices=torch.arange(nv, dtype=torch.int32, device="cuda"))
    labels = g.renumber_map().cpu().numpy()
    offsets, indices, _ = cpu_oracle.csc_from_coo(s.cpu().numpy(), d.cpu().numpy(), None, nv)
    return g, labels, offsets, indices, nv


def _report(tag, ext, ref32):
    err = util.rel_err(ext, ref32)
    overlap = len(set(util.top_k(ext, 100)) & set(util.top_k(ref32, 100)))
    print(f"\n[{tag}] vs fp32 pagerank_reference: max rel err {err.max():.3e}, "
          f"top-100 overlap {overlap}/100")


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_pagerank_rmat_s20_vs_reference(handle, dtype):
    """BASELINE configs[1]: R-MAT scale 20 edge factor 16, to convergence (alpha 0.85, 1e-6)."""
    from cugraph_forked_b200 import engine
    from oracle import cpu_oracle
    g, labels, offsets, indices, nv = _config(handle, 20)
    pr, st = engine.pagerank(g, dtype=TDT[dtype], alpha=0.85, epsilon=1e-6, max_iterations=500)
    assert st.converged and st.gpu_launches >= st.iterations
    ext = util.to_external(pr.cpu().numpy(), labels, nv)
    assert abs(float(ext.astype(np.float64).sum()) - 1.0) < (1e-4 if dtype == np.float32 else 1e-9)
    # (1) the tolerance north_star states, at the GPU's iteration count.  fp32: against the fp64
    #     port.  fp64: against the port with its sums in extended precision - the reference's
    #     sequential fp64 row sums carry up to in-degree * 2^-53 of rounding error themselves
    #     (4e-12 on this graph's hub rows, measured on B200 in call r2b), so the plain port is
    #     held to that bound and the extended one to 1e-12; all three differences are printed.
    ref64, _ = cpu_oracle.pagerank(offsets, indices, None, dtype=np.float64, alpha=0.85,
                                   fixed_iterations=st.iterations)
    err = util.rel_err(ext, ref64)
    if dtype == np.float64:
        refx, _ = cpu_oracle.pagerank(offsets, indices, None, dtype=np.float64, alpha=0.85,
                                      fixed_iterations=st.iterations, impl="ext")
        errx = util.rel_err(ext, refx)
        max_deg = int(np.diff(offsets).max())
        print(f"\n[s20 f64] max rel err: GPU vs extended-sum oracle {errx.max():.3e}, GPU vs "
              f"pagerank_reference port {err.max():.3e}, port vs extended {util.rel_err(ref64, refx).max():.3e}"
              f" (max in-degree {max_deg})")
        assert errx.max() <= 1e-12, f"max rel err vs extended-sum oracle {errx.max():.3e}"
        assert err.max() <= 1e-12 + max_deg * 2.0**-53
        assert _same_top(ext, refx)
    else:
        assert err.max() <= TOL[dtype], f"max rel err vs fp64 oracle {err.max():.3e}"
    assert _same_top(ext, ref64)
    # (2) the reference's own compiled pagerank_reference in fp64, to convergence, at the
    #     reference's own bound (pagerank_test.cpp:355-362)
    impl = "ref" if cpu_oracle.ref_available() else "port"
    ref, _ = cpu_oracle.pagerank(offsets, indices, None, dtype=np.float64, alpha=0.85,
                                 epsilon=1e-6, max_iterations=500, impl=impl)
    thr = np.maximum(np.maximum(ext, ref) * 1e-3, 1e-3 / nv)
    assert np.all(np.abs(ext - ref) < thr)
    if dtype == np.float64:
        # fp64: the reference itself is inside the tolerance, and stops at the same iteration
        _, it_port = cpu_oracle.pagerank(offsets, indices, None, dtype=np.float64, alpha=0.85,
                                         epsilon=1e-6, max_iterations=500)
        assert it_port == st.iterations
        assert util.rel_err(ext, ref).max() <= 1e-12 + max_deg * 2.0**-53
        assert _same_top(ext, ref)
    else:
        # fp32 pagerank_reference sums every row sequentially in fp32: on the hub rows of this
        # graph (in-degree > 1e5) it is itself ~3e-3 away from its own fp64 result (measured on
        # B200, call r2b: 3.1571e-3 vs 3.1472e-3 for the top vertex), i.e. outside the reference's
        # own 1e-3 bound.  Reported, not asserted.
        ref32, _ = cpu_oracle.pagerank(offsets, indices, None, dtype=np.float32, alpha=0.85,
                                       epsilon=1e-6, max_iterations=500, impl=impl)
        _report("s20 f32", ext, ref32)
        e32 = util.rel_err(ref32, ref)
        print(f"[s20 f32] fp32 pagerank_reference vs its own fp64 run: max rel err {e32.max():.3e}")


def test_pagerank_rmat_s24_f32_vs_reference(handle):
    """BASELINE configs[2] (the headline bench workload): R-MAT scale 24, 5 pinned iterations of
    the reference's pagerank_reference (fp64 for the tolerance, fp32 reported)."""
    from cugraph_forked_b200 import engine
    from oracle import cpu_oracle
    if torch.cuda.mem_get_info()[0] < 40 * 2**30:
        pytest.skip("needs ~30 GB of free HBM")
    iters = 5
    g, labels, offsets, indices, nv = _config(handle, 24)
    pr, st = engine.pagerank(g, dtype=torch.float32, alpha=0.85, fixed_iterations=iters)
    assert st.iterations == iters
    ext = util.to_external(pr.cpu().numpy(), labels, nv)
    impl = "ref" if cpu_oracle.ref_available() else "port"
    ref64, _ = cpu_oracle.pagerank(offsets, indices, None, dtype=np.float64, alpha=0.85,
                                   fixed_iterations=iters, impl=impl)
    err = util.rel_err(ext, ref64)
    assert err.max() <= 1e-5, f"max rel err vs fp64 pagerank_reference {err.max():.3e}"
    assert _same_top(ext, ref64)
    ref32, _ = cpu_oracle.pagerank(offsets, indices, None, dtype=np.float32, alpha=0.85,
                                   fixed_iterations=iters, impl=impl)
    _report("s24 f32, 5 iterations", ext, ref32)    # reported only: see the s20 case
