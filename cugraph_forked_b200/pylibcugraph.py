"""pylibcugraph-shaped host layer over the cugraph_c C ABI of libcugraph_b200.so.

Mirrors the reference's Python binding for the pull path, same names and argument meaning:
  ResourceHandle          python/pylibcugraph/pylibcugraph/resource_handle.pyx:30
  GraphProperties         python/pylibcugraph/pylibcugraph/graph_properties.pyx:21
  SGGraph / MGGraph       python/pylibcugraph/pylibcugraph/graphs.pyx:130,300
  pagerank                python/pylibcugraph/pylibcugraph/pagerank.pyx:57
  personalized_pagerank   python/pylibcugraph/pylibcugraph/personalized_pagerank.pyx:57
  katz_centrality         python/pylibcugraph/pylibcugraph/katz_centrality.pyx:57
  eigenvector_centrality  python/pylibcugraph/pylibcugraph/eigenvector_centrality.pyx:57
Error codes map to Python exceptions as utils.pyx:37-82 does.  Inputs are anything with
`__cuda_array_interface__` (torch CUDA tensors, cupy arrays); results come back as torch CUDA
tensors (vertices, values), copied out of the C result object exactly as the reference copies
into cupy arrays (utils.pyx copy_to_cupy_array).

Every call goes to the CUDA library; there is no CPU path here.
"""
from __future__ import annotations

import ctypes
from ctypes import POINTER, byref, c_char_p, c_double, c_int, c_size_t, c_void_p

import torch

from . import _lib

INT32, INT64, FLOAT32, FLOAT64 = 0, 1, 2, 3
_TYPESTR_TO_ID = {"<i4": INT32, "<i8": INT64, "<f4": FLOAT32, "<f8": FLOAT64}
_ID_TO_TORCH = {INT32: torch.int32, INT64: torch.int64, FLOAT32: torch.float32,
                FLOAT64: torch.float64}

_EXC = {
    1: RuntimeError,          # CUGRAPH_UNKNOWN_ERROR
    2: ValueError,            # CUGRAPH_INVALID_HANDLE
    3: MemoryError,           # CUGRAPH_ALLOC_ERROR
    4: ValueError,            # CUGRAPH_INVALID_INPUT
    5: NotImplementedError,   # CUGRAPH_NOT_IMPLEMENTED
    6: ValueError,            # CUGRAPH_UNSUPPORTED_TYPE_COMBINATION
}

_P = c_void_p
_PP = POINTER(c_void_p)

# name -> (restype, argtypes) for the cugraph_c entry points this module binds
_C_SIGNATURES = {
    "cugraph_error_message": (c_char_p, [_P]),
    "cugraph_error_free": (None, [_P]),
    "cugraph_create_resource_handle": (_P, [_P]),
    "cugraph_resource_handle_get_rank": (c_int, [_P]),
    "cugraph_free_resource_handle": (None, [_P]),
    "cugraph_type_erased_device_array_view_create": (_P, [_P, c_size_t, c_int]),
    "cugraph_type_erased_device_array_view_free": (None, [_P]),
    "cugraph_type_erased_device_array_view_size": (c_size_t, [_P]),
    "cugraph_type_erased_device_array_view_type": (c_int, [_P]),
    "cugraph_type_erased_device_array_view_pointer": (_P, [_P]),
    "cugraph_sg_graph_create": (c_int, [_P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, _PP, _PP]),
    "cugraph_sg_graph_free": (None, [_P]),
    "cugraph_mg_graph_create": (c_int, [_P, _P, _P, _P, _P, _P, _P, c_int, c_size_t, c_int, _PP,
                                        _PP]),
    "cugraph_mg_graph_free": (None, [_P]),
    "cugraph_pagerank": (c_int, [_P, _P, _P, _P, _P, _P, c_double, c_double, c_size_t, c_int,
                                 _PP, _PP]),
    "cugraph_personalized_pagerank": (c_int, [_P, _P, _P, _P, _P, _P, _P, _P, c_double, c_double,
                                              c_size_t, c_int, _PP, _PP]),
    "cugraph_katz_centrality": (c_int, [_P, _P, _P, c_double, c_double, c_double, c_size_t, c_int,
                                        _PP, _PP]),
    "cugraph_eigenvector_centrality": (c_int, [_P, _P, c_double, c_size_t, c_int, _PP, _PP]),
    "cugraph_hits": (c_int, [_P, _P, c_double, c_size_t, _P, _P, c_int, c_int, _PP, _PP]),
    "cugraph_hits_result_get_vertices": (_P, [_P]),
    "cugraph_hits_result_get_hubs": (_P, [_P]),
    "cugraph_hits_result_get_authorities": (_P, [_P]),
    "cugraph_hits_result_get_hub_score_differences": (c_double, [_P]),
    "cugraph_hits_result_get_number_of_iterations": (c_size_t, [_P]),
    "cugraph_hits_result_free": (None, [_P]),
    "cugraph_centrality_result_get_vertices": (_P, [_P]),
    "cugraph_centrality_result_get_values": (_P, [_P]),
    "cugraph_centrality_result_free": (None, [_P]),
}

_bound = None


def _c():
    global _bound
    if _bound is None:
        lib = _lib.lib()
        for name, (res, args) in _C_SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _bound = lib
    return _bound


def assert_success(code: int, err: c_void_p, api_name: str) -> None:
    """utils.pyx:37-82 assert_success: raise the exception class the reference raises."""
    if code == 0:
        return
    msg = ""
    if err:
        raw = _c().cugraph_error_message(err)
        msg = raw.decode() if raw else ""
        _c().cugraph_error_free(err)
    name = _lib.ERROR_NAMES.get(code, "unknown error code")
    raise _EXC.get(code, RuntimeError)(
        f"non-success value returned from {api_name}: {name} {msg}")


def _assert_cai(obj, var_name: str, allow_none: bool = False) -> None:
    """utils.pyx:85-95 assert_CAI_type."""
    if obj is None and allow_none:
        return
    if not hasattr(obj, "__cuda_array_interface__"):
        raise TypeError(f"{var_name} must be None or support __cuda_array_interface__"
                        if allow_none else
                        f"{var_name} does not support __cuda_array_interface__")


class _View:
    """Owns one cugraph_type_erased_device_array_view_t over caller memory."""

    def __init__(self, obj):
        self.ptr = None
        if obj is None:
            return
        cai = obj.__cuda_array_interface__
        if cai["typestr"] not in _TYPESTR_TO_ID:
            raise TypeError(f"unsupported dtype {cai['typestr']}")
        if cai.get("strides") is not None and len(cai["shape"]) == 1:
            itemsize = int(cai["typestr"][2:])
            if cai["strides"][0] != itemsize:
                raise ValueError("array must be contiguous")
        n = 1
        for s in cai["shape"]:
            n *= s
        self.keepalive = obj
        self.ptr = _c().cugraph_type_erased_device_array_view_create(
            cai["data"][0], n, _TYPESTR_TO_ID[cai["typestr"]])

    def __del__(self):
        if getattr(self, "ptr", None):
            _c().cugraph_type_erased_device_array_view_free(self.ptr)
            self.ptr = None


def _copy_out(view_ptr) -> torch.Tensor:
    """Copy a result view into a fresh torch tensor (device to device) and free the view."""
    c = _c()
    n = c.cugraph_type_erased_device_array_view_size(view_ptr)
    dt = _ID_TO_TORCH[c.cugraph_type_erased_device_array_view_type(view_ptr)]
    out = torch.empty(n, dtype=dt, device="cuda")
    if n:
        src = c.cugraph_type_erased_device_array_view_pointer(view_ptr)

        class _Raw:
            pass

        raw = _Raw()
        raw.__cuda_array_interface__ = {
            "shape": (n,), "typestr": {v: k for k, v in _TYPESTR_TO_ID.items()}[
                c.cugraph_type_erased_device_array_view_type(view_ptr)],
            "data": (src, False), "version": 2}
        out.copy_(torch.as_tensor(raw, device="cuda"))
        torch.cuda.current_stream().synchronize()
    c.cugraph_type_erased_device_array_view_free(view_ptr)
    return out


class ResourceHandle:
    """resource_handle.pyx:30.  `handle` is None (library-owned stream) or the integer address
    of a b200_handle_t made with b200_handle_create / b200_handle_create_mg, the stand-in for
    the raft::handle_t address the reference takes."""

    def __init__(self, handle=None):
        self._engine_handle = None
        if handle is not None and not isinstance(handle, int):
            # an engine.Handle (SG or multi_gpu): keep it alive and pass its address
            self._engine_handle = handle
            handle = handle._h.value
        self.c_resource_handle_ptr = _c().cugraph_create_resource_handle(handle)
        if not self.c_resource_handle_ptr:
            raise RuntimeError("cugraph_create_resource_handle failed: "
                               + (_lib.lib().b200_last_error() or b"").decode())

    def get_rank(self) -> int:
        return _c().cugraph_resource_handle_get_rank(self.c_resource_handle_ptr)

    def __del__(self):
        if getattr(self, "c_resource_handle_ptr", None):
            _c().cugraph_free_resource_handle(self.c_resource_handle_ptr)
            self.c_resource_handle_ptr = None


class _GraphPropertiesStruct(ctypes.Structure):
    _fields_ = [("is_symmetric", c_int), ("is_multigraph", c_int)]


class GraphProperties:
    """graph_properties.pyx:21."""

    def __init__(self, is_symmetric: bool = False, is_multigraph: bool = False):
        self.c_graph_properties = _GraphPropertiesStruct(int(is_symmetric), int(is_multigraph))

    @property
    def is_symmetric(self) -> bool:
        return bool(self.c_graph_properties.is_symmetric)

    @is_symmetric.setter
    def is_symmetric(self, value) -> None:
        self.c_graph_properties.is_symmetric = int(bool(value))

    @property
    def is_multigraph(self) -> bool:
        return bool(self.c_graph_properties.is_multigraph)

    @is_multigraph.setter
    def is_multigraph(self, value) -> None:
        self.c_graph_properties.is_multigraph = int(bool(value))


class _GPUGraph:
    c_graph_ptr = None


class SGGraph(_GPUGraph):
    """graphs.pyx:130: single-GPU graph from a device COO."""

    def __init__(self, resource_handle, graph_properties, src_array, dst_array, weight_array,
                 store_transposed=False, renumber=False, do_expensive_check=False,
                 edge_id_array=None, edge_type_array=None):
        _assert_cai(src_array, "src_array")
        _assert_cai(dst_array, "dst_array")
        _assert_cai(weight_array, "weight_array", allow_none=True)
        _assert_cai(edge_id_array, "edge_id_array", allow_none=True)
        _assert_cai(edge_type_array, "edge_type_array", allow_none=True)
        self._handle = resource_handle
        views = [_View(a) for a in (src_array, dst_array, weight_array, edge_id_array,
                                    edge_type_array)]
        g, err = c_void_p(), c_void_p()
        _sync_inputs()
        code = _c().cugraph_sg_graph_create(
            resource_handle.c_resource_handle_ptr, byref(graph_properties.c_graph_properties),
            views[0].ptr, views[1].ptr, views[2].ptr, views[3].ptr, views[4].ptr,
            int(store_transposed), int(renumber), int(do_expensive_check), byref(g), byref(err))
        assert_success(code, err, "cugraph_sg_graph_create()")
        self.c_graph_ptr = g

    def __del__(self):
        if getattr(self, "c_graph_ptr", None):
            _c().cugraph_sg_graph_free(self.c_graph_ptr)
            self.c_graph_ptr = None


class MGGraph(_GPUGraph):
    """graphs.pyx:300: this rank's slice of a distributed COO; edges are shuffled to their 2D
    owners and renumbered inside the library."""

    def __init__(self, resource_handle, graph_properties, src_array, dst_array, weight_array,
                 store_transposed=False, num_edges=-1, do_expensive_check=False,
                 edge_id_array=None, edge_type_array=None):
        _assert_cai(src_array, "src_array")
        _assert_cai(dst_array, "dst_array")
        _assert_cai(weight_array, "weight_array", allow_none=True)
        self._handle = resource_handle
        views = [_View(a) for a in (src_array, dst_array, weight_array, edge_id_array,
                                    edge_type_array)]
        g, err = c_void_p(), c_void_p()
        _sync_inputs()
        code = _c().cugraph_mg_graph_create(
            resource_handle.c_resource_handle_ptr, byref(graph_properties.c_graph_properties),
            views[0].ptr, views[1].ptr, views[2].ptr, views[3].ptr, views[4].ptr,
            int(store_transposed), max(int(num_edges), 0), int(do_expensive_check), byref(g),
            byref(err))
        assert_success(code, err, "cugraph_mg_graph_create()")
        self.c_graph_ptr = g

    def __del__(self):
        if getattr(self, "c_graph_ptr", None):
            _c().cugraph_mg_graph_free(self.c_graph_ptr)
            self.c_graph_ptr = None


def _sync_inputs() -> None:
    # the library works on its own stream: inputs produced on torch's stream must be complete
    torch.cuda.current_stream().synchronize()


def _finish(result: c_void_p):
    c = _c()
    vertices = _copy_out(c.cugraph_centrality_result_get_vertices(result))
    values = _copy_out(c.cugraph_centrality_result_get_values(result))
    c.cugraph_centrality_result_free(result)
    return vertices, values


def pagerank(resource_handle, graph, precomputed_vertex_out_weight_vertices,
             precomputed_vertex_out_weight_sums, initial_guess_vertices, initial_guess_values,
             alpha, epsilon, max_iterations, do_expensive_check):
    """pagerank.pyx:57.  Returns (vertices, values)."""
    _assert_cai(precomputed_vertex_out_weight_vertices,
                "precomputed_vertex_out_weight_vertices", allow_none=True)
    _assert_cai(precomputed_vertex_out_weight_sums, "precomputed_vertex_out_weight_sums",
                allow_none=True)
    _assert_cai(initial_guess_vertices, "initial_guess_vertices", allow_none=True)
    _assert_cai(initial_guess_values, "initial_guess_values", allow_none=True)
    v = [_View(a) for a in (precomputed_vertex_out_weight_vertices,
                            precomputed_vertex_out_weight_sums, initial_guess_vertices,
                            initial_guess_values)]
    res, err = c_void_p(), c_void_p()
    _sync_inputs()
    code = _c().cugraph_pagerank(resource_handle.c_resource_handle_ptr, graph.c_graph_ptr,
                                 v[0].ptr, v[1].ptr, v[2].ptr, v[3].ptr, alpha, epsilon,
                                 max_iterations, int(do_expensive_check), byref(res), byref(err))
    assert_success(code, err, "cugraph_pagerank")
    return _finish(res)


def personalized_pagerank(resource_handle, graph, precomputed_vertex_out_weight_vertices,
                          precomputed_vertex_out_weight_sums, initial_guess_vertices,
                          initial_guess_values, personalization_vertices, personalization_values,
                          alpha, epsilon, max_iterations, do_expensive_check):
    """personalized_pagerank.pyx:57.  Returns (vertices, values)."""
    _assert_cai(personalization_vertices, "personalization_vertices", allow_none=True)
    _assert_cai(personalization_values, "personalization_values", allow_none=True)
    v = [_View(a) for a in (precomputed_vertex_out_weight_vertices,
                            precomputed_vertex_out_weight_sums, initial_guess_vertices,
                            initial_guess_values, personalization_vertices,
                            personalization_values)]
    res, err = c_void_p(), c_void_p()
    _sync_inputs()
    code = _c().cugraph_personalized_pagerank(
        resource_handle.c_resource_handle_ptr, graph.c_graph_ptr, v[0].ptr, v[1].ptr, v[2].ptr,
        v[3].ptr, v[4].ptr, v[5].ptr, alpha, epsilon, max_iterations, int(do_expensive_check),
        byref(res), byref(err))
    assert_success(code, err, "cugraph_personalized_pagerank")
    return _finish(res)


def katz_centrality(resource_handle, graph, betas, alpha, beta, epsilon, max_iterations,
                    do_expensive_check):
    """katz_centrality.pyx:57.  Returns (vertices, values).  As in the reference C API the
    per-vertex `betas` are accepted but not used (c_api/katz.cpp:158)."""
    _assert_cai(betas, "betas", allow_none=True)
    b = _View(betas)
    res, err = c_void_p(), c_void_p()
    _sync_inputs()
    code = _c().cugraph_katz_centrality(resource_handle.c_resource_handle_ptr, graph.c_graph_ptr,
                                        b.ptr, alpha, beta, epsilon, max_iterations,
                                        int(do_expensive_check), byref(res), byref(err))
    assert_success(code, err, "cugraph_katz_centrality")
    return _finish(res)


def eigenvector_centrality(resource_handle, graph, epsilon, max_iterations, do_expensive_check):
    """eigenvector_centrality.pyx:57.  Returns (vertices, values)."""
    res, err = c_void_p(), c_void_p()
    _sync_inputs()
    code = _c().cugraph_eigenvector_centrality(
        resource_handle.c_resource_handle_ptr, graph.c_graph_ptr, epsilon, max_iterations,
        int(do_expensive_check), byref(res), byref(err))
    assert_success(code, err, "cugraph_eigenvector_centrality")
    return _finish(res)


def hits(resource_handle, graph, tol, max_iter, initial_hubs_guess_vertices,
         initial_hubs_guess_values, normalized, do_expensive_check):
    """hits.pyx:58.  Returns (vertices, hubs, authorities)."""
    _assert_cai(initial_hubs_guess_vertices, "initial_hubs_guess_vertices", allow_none=True)
    _assert_cai(initial_hubs_guess_values, "initial_hubs_guess_values", allow_none=True)
    v = [_View(a) for a in (initial_hubs_guess_vertices, initial_hubs_guess_values)]
    res, err = c_void_p(), c_void_p()
    _sync_inputs()
    c = _c()
    code = c.cugraph_hits(resource_handle.c_resource_handle_ptr, graph.c_graph_ptr, tol, max_iter,
                          v[0].ptr, v[1].ptr, int(normalized), int(do_expensive_check), byref(res),
                          byref(err))
    assert_success(code, err, "cugraph_hits")
    vertices = _copy_out(c.cugraph_hits_result_get_vertices(res))
    hubs = _copy_out(c.cugraph_hits_result_get_hubs(res))
    authorities = _copy_out(c.cugraph_hits_result_get_authorities(res))
    c.cugraph_hits_result_free(res)
    return vertices, hubs, authorities
