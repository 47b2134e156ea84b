"""GPU parity of the two-phase pull path - the round-2 default "lane owns a segment" layout
(csrc/sell.cu[h], B200_LAYOUT=2) and the round-1 banded layout kept as a fallback
(csrc/banded.cu[h], B200_LAYOUT=1): the same
oracle comparisons as test_gpu_centrality.py with B200_BANDED=1, plus agreement with the
single-phase kernel and a torch fp64 segmented sum.  `st.gpu_launches >= 3 * iterations`
(band + tiles + tail per iteration) proves the banded kernels actually ran."""
import numpy as np
import pytest
import torch

from tests import util
from tests.test_gpu_centrality import TDT, _check_pagerank, _graph

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["2", "1"], ids=["sell", "banded-r1"])
def banded_on(monkeypatch, request):
    monkeypatch.setenv("B200_BANDED", "1")
    monkeypatch.setenv("B200_LAYOUT", request.param)


def _dense_ref(g, x, weighted):
    nv = g.num_vertices
    offs = g.offsets().long()
    idx = g.indices().long()
    rows = torch.repeat_interleave(torch.arange(nv, device="cuda"), offs[1:] - offs[:-1])
    vals = x.double()[idx] * (g.weights().double() if weighted else 1.0)
    return torch.zeros(nv, device="cuda", dtype=torch.float64).index_add_(0, rows, vals)


@pytest.mark.parametrize("scale,hot_min", [(13, 1), (16, 32768), (18, 32768), (18, 1)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("weighted", [False, True])
def test_spmv_banded(handle, monkeypatch, scale, hot_min, dtype, weighted):
    from cugraph_forked_b200 import engine
    monkeypatch.setenv("B200_BAND_HOT_MIN", str(hot_min))
    src, dst = util.rmat(scale, seed=9)
    w = np.random.default_rng(2).random(len(src)).astype(np.float32) if weighted else None
    g, _ = _graph(handle, src, dst, w, 1 << scale)
    nv = g.num_vertices
    x = torch.rand(nv, device="cuda", dtype=dtype)
    y = engine.spmv_pull(g, x, 2.0, 0.5)
    ref = _dense_ref(g, x, weighted) * 2.0 + 0.5
    tol = 2e-6 if dtype == torch.float32 else 1e-13
    assert torch.max(torch.abs(y.double() - ref) / ref.abs().clamp_min(1e-30)).item() <= tol
    # run-to-run determinism, and agreement with the single-phase kernel
    assert torch.equal(y, engine.spmv_pull(g, x, 2.0, 0.5))
    monkeypatch.setenv("B200_BANDED", "0")
    y0 = engine.spmv_pull(g, x, 2.0, 0.5)
    assert torch.allclose(y.double(), y0.double(), rtol=4e-6 if dtype == torch.float32 else 1e-13,
                          atol=0)


@pytest.mark.parametrize("scale", [14, 18])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_pagerank_banded(handle, scale, dtype):
    src, dst = util.rmat(scale)
    st = _check_pagerank(handle, src, dst, None, 1 << scale, dtype)
    assert st.gpu_launches >= 3 * st.iterations


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_pagerank_banded_weighted_personalized(handle, monkeypatch, dtype):
    monkeypatch.setenv("B200_BAND_HOT_MIN", "1")
    scale = 13
    src, dst = util.rmat(scale, seed=3)
    nv = 1 << scale
    rng = np.random.default_rng(11)
    w = rng.random(len(src)).astype(dtype)
    st = _check_pagerank(handle, src, dst, w, nv, dtype)
    assert st.gpu_launches >= 3 * st.iterations
    pv = rng.permutation(nv)[: nv // 2].astype(np.int32)
    pw = rng.random(nv // 2).astype(dtype)
    _check_pagerank(handle, src, dst, w, nv, dtype, personalization=(pv, pw))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_katz_banded(handle, dtype):
    from cugraph_forked_b200 import engine
    scale = 14
    src, dst = util.rmat(scale, seed=2)
    nv = 1 << scale
    w = np.random.default_rng(3).random(len(src)).astype(dtype)
    g, labels = _graph(handle, src, dst, w, nv)
    alpha = 0.5 / max(1.0, float(np.bincount(dst).max()))
    kc, st = engine.katz_centrality(g, alpha=alpha, beta=1.0, dtype=TDT[dtype], normalize=True)
    assert st.gpu_launches >= 3 * st.iterations
    ref, _ = util.oracle_katz(src, dst, w, nv, dtype=np.float64, alpha=alpha, beta=1.0,
                              normalize=True, fixed_iterations=st.iterations)
    err = util.rel_err(util.to_external(kc.cpu().numpy(), labels, nv), ref)
    assert err.max() <= (1e-5 if dtype == np.float32 else 1e-11)


def test_pagerank_banded_scale20_properties(handle):
    """BASELINE.json configs[1] size: banded fp32 / fp64 against the single-phase fp64 result."""
    import os
    from cugraph_forked_b200 import engine
    scale = 20
    nv, ne = 1 << scale, 16 << scale
    s, d = engine.generate_rmat_edgelist(handle, scale, ne, seed=0)
    g = engine.Graph(handle, s, d, vertices=torch.arange(nv, dtype=torch.int32, device="cuda"))
    p64, st64 = engine.pagerank(g, dtype=torch.float64)
    p32, st32 = engine.pagerank(g, dtype=torch.float32)
    assert st64.gpu_launches >= 3 * st64.iterations and st32.gpu_launches >= 3 * st32.iterations
    os.environ["B200_BANDED"] = "0"
    q64, su64 = engine.pagerank(g, dtype=torch.float64)
    assert su64.iterations == st64.iterations
    assert ((p64 - q64).abs() / q64).max().item() < 1e-12
    assert abs(p64.sum().item() - 1.0) < 1e-9
    assert ((p32.double() - q64).abs() / q64).max().item() < 1e-5
    assert torch.equal(torch.topk(q64, 100).indices, torch.topk(p32.double(), 100).indices)
