"""GPU: b200_spmv_pull_noalloc (csrc/noalloc.cu) - the pull SpMV for callers that may not
allocate; its scratch is a static __device__ array.  No counterpart in the reference,
so the checker is the existing entry point b200_spmv_pull on the
same graph and vector, within the tolerances tests/test_gpu_centrality.py uses for the raw SpMV
(2e-6 relative in fp32, 1e-13 in fp64), on two calls in a row (the second call meets the first
call's leftovers in the shared scratch array).  Collected last
(pytest --collect-only)."""
import ctypes

import numpy as np
import pytest
import torch

from tests import util

pytestmark = pytest.mark.gpu


def _value_type(dtype):
    from cugraph_forked_b200 import engine
    return engine._DT[dtype]


def _noalloc(g, x, scale, init):
    """through the C ABI; asynchronous on the handle's stream (a torch side stream here)"""
    from cugraph_forked_b200 import _lib, engine
    y = torch.full_like(x, float("nan"))
    torch.cuda.current_stream().synchronize()     # x and y are ready before the side stream runs
    before = torch.cuda.memory_allocated()
    engine.check(_lib.lib().b200_spmv_pull_noalloc(
        g.handle._h, g._g, _value_type(x.dtype), ctypes.c_void_p(x.data_ptr()),
        ctypes.c_double(scale), ctypes.c_double(init), ctypes.c_void_p(y.data_ptr())))
    assert torch.cuda.memory_allocated() == before
    g.handle.stream.synchronize()
    return y


@pytest.fixture(scope="module")
def stream_handle():
    from cugraph_forked_b200 import engine
    # a real side stream: torch's default stream has the handle value 0, which the library reads
    # as "create a stream of your own"
    h = engine.Handle(stream=torch.cuda.Stream())
    yield h
    h.close()


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("weighted", [False, True])
def test_noalloc_matches_spmv_pull_twice(stream_handle, dtype, weighted):
    from cugraph_forked_b200 import engine
    scale = 16   # 1 M edges = 256 windows; hub rows of an R-MAT graph span many of them
    src, dst = util.rmat(scale, seed=9)
    nv = 1 << scale
    w = np.random.default_rng(2).random(len(src)).astype(np.float32) if weighted else None
    g = engine.Graph(stream_handle, torch.from_numpy(src), torch.from_numpy(dst),
                     None if w is None else torch.from_numpy(w),
                     vertices=torch.arange(nv, dtype=torch.int32))
    tol = 2e-6 if dtype == torch.float32 else 1e-13
    gen = torch.Generator(device="cuda").manual_seed(5)
    first = None
    for call in range(2):
        x = torch.rand(nv, device="cuda", dtype=dtype, generator=gen) + 0.5
        ref = engine.spmv_pull(g, x, 2.0, 0.5)        # the existing entry point
        got = _noalloc(g, x, 2.0, 0.5)
        assert not torch.isnan(got).any()
        rel = ((got.double() - ref.double()).abs() / ref.double().abs().clamp_min(1e-30)).max().item()
        assert rel <= tol, f"call {call}: max rel diff {rel:.3e}"
        again = _noalloc(g, x, 2.0, 0.5)              # fixed order: bit-identical repeat
        assert torch.equal(got, again)
        if first is None:
            first = got
    # rows without edges get `init`
    deg = g.offsets().long()
    empty = (deg[1:] - deg[:-1]) == 0
    if empty.any():
        assert torch.all(got[empty] == 0.5)


def test_noalloc_rejects_bad_arguments(stream_handle):
    from cugraph_forked_b200 import _lib, engine
    src, dst = util.rmat(8)
    g = engine.Graph(stream_handle, torch.from_numpy(src), torch.from_numpy(dst))
    x = torch.ones(g.num_vertices, device="cuda")
    rc = _lib.lib().b200_spmv_pull_noalloc(g.handle._h, g._g, 7, ctypes.c_void_p(x.data_ptr()),
                                           ctypes.c_double(1.0), ctypes.c_double(0.0),
                                           ctypes.c_void_p(x.data_ptr()))
    assert rc != 0 and b"value_type" in _lib.lib().b200_last_error()
    rc = _lib.lib().b200_spmv_pull_noalloc(g.handle._h, g._g, 0, None, ctypes.c_double(1.0),
                                           ctypes.c_double(0.0), ctypes.c_void_p(x.data_ptr()))
    assert rc != 0
