"""ctypes binding of libcugraph_b200.so (include/b200_cugraph.h).

The library is the product: there is no CPU or PyTorch fallback.  Loading fails loudly when the
shared object is missing (build it with `python -c "import __graft_entry__ as g; g.build()"` or
`make -C cugraph_forked_b200/csrc`).
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200_LIBRARY points tuning sweeps at a variant build (scripts/build_variant.sh); it is still
# this library, compiled with other constants - never a fallback
LIB_PATH = os.environ.get("B200_LIBRARY") or os.path.join(_HERE, "libcugraph_b200.so")

INT32, INT64, FLOAT32, FLOAT64 = 0, 1, 2, 3

# cugraph_error_code_t (cpp/include/cugraph_c/error.h:25-33)
ERROR_NAMES = {
    0: "CUGRAPH_SUCCESS",
    1: "CUGRAPH_UNKNOWN_ERROR",
    2: "CUGRAPH_INVALID_HANDLE",
    3: "CUGRAPH_ALLOC_ERROR",
    4: "CUGRAPH_INVALID_INPUT",
    5: "CUGRAPH_NOT_IMPLEMENTED",
    6: "CUGRAPH_UNSUPPORTED_TYPE_COMBINATION",
}


class GraphView(ctypes.Structure):
    _fields_ = [
        ("num_vertices", ctypes.c_int64),
        ("num_edges", ctypes.c_int64),
        ("vertex_type", ctypes.c_int32),
        ("edge_type", ctypes.c_int32),
        ("weight_type", ctypes.c_int32),
        ("store_transposed", ctypes.c_int32),
        ("renumbered", ctypes.c_int32),
        ("num_segments", ctypes.c_int32),
        ("segment_offsets", ctypes.c_int64 * 8),
        ("offsets", ctypes.c_void_p),
        ("indices", ctypes.c_void_p),
        ("weights", ctypes.c_void_p),
        ("renumber_map", ctypes.c_void_p),
        ("comm_rank", ctypes.c_int32),
        ("comm_size", ctypes.c_int32),
        ("row_comm_size", ctypes.c_int32),
        ("col_comm_size", ctypes.c_int32),
        ("local_vertex_first", ctypes.c_int64),
        ("local_vertex_last", ctypes.c_int64),
        ("major_first", ctypes.c_int64),
        ("major_last", ctypes.c_int64),
        ("minor_first", ctypes.c_int64),
        ("minor_last", ctypes.c_int64),
        ("dcs_offsets", ctypes.c_void_p),
        ("dcs_nzd_vertices", ctypes.c_void_p),
        ("dcs_nzd_count", ctypes.c_int64),
        ("major_hypersparse_first", ctypes.c_int64),
    ]


class CentralityStats(ctypes.Structure):
    _fields_ = [
        ("iterations", ctypes.c_int64),
        ("converged", ctypes.c_int32),
        ("last_diff", ctypes.c_double),
        ("elapsed_ms", ctypes.c_double),
        ("gpu_launches", ctypes.c_int64),
        ("spmv_ms", ctypes.c_double),
        ("spmv_launches", ctypes.c_int64),
    ]


class PagerankParams(ctypes.Structure):
    _fields_ = [
        ("alpha", ctypes.c_double),
        ("epsilon", ctypes.c_double),
        ("max_iterations", ctypes.c_uint64),
        ("value_type", ctypes.c_int32),
        ("has_initial_guess", ctypes.c_int32),
        ("precomputed_out_weight_sums", ctypes.c_void_p),
        ("personalization_vertices", ctypes.c_void_p),
        ("personalization_values", ctypes.c_void_p),
        ("personalization_size", ctypes.c_uint64),
        ("fixed_iterations", ctypes.c_int64),
    ]


class KatzParams(ctypes.Structure):
    _fields_ = [
        ("alpha", ctypes.c_double),
        ("beta", ctypes.c_double),
        ("epsilon", ctypes.c_double),
        ("max_iterations", ctypes.c_uint64),
        ("value_type", ctypes.c_int32),
        ("has_initial_guess", ctypes.c_int32),
        ("normalize", ctypes.c_int32),
        ("betas", ctypes.c_void_p),
        ("fixed_iterations", ctypes.c_int64),
    ]


class EigenvectorParams(ctypes.Structure):
    _fields_ = [
        ("epsilon", ctypes.c_double),
        ("max_iterations", ctypes.c_uint64),
        ("value_type", ctypes.c_int32),
        ("has_initial_guess", ctypes.c_int32),
        ("fixed_iterations", ctypes.c_int64),
    ]


class HitsParams(ctypes.Structure):
    _fields_ = [
        ("epsilon", ctypes.c_double),
        ("max_iterations", ctypes.c_uint64),
        ("value_type", ctypes.c_int32),
        ("has_initial_hubs_guess", ctypes.c_int32),
        ("normalize", ctypes.c_int32),
        ("pad", ctypes.c_int32),
        ("fixed_iterations", ctypes.c_int64),
    ]


class HostEdgelistInfo(ctypes.Structure):
    _fields_ = [
        ("num_rows", ctypes.c_int64),
        ("num_cols", ctypes.c_int64),
        ("num_edges", ctypes.c_int64),
        ("file_entries", ctypes.c_int64),
        ("is_symmetric", ctypes.c_int32),
        ("is_pattern", ctypes.c_int32),
        ("vertex_type", ctypes.c_int32),
        ("weight_type", ctypes.c_int32),
        ("pinned", ctypes.c_int32),
        ("parser", ctypes.c_int32),
    ]


# every symbol include/b200_cugraph.h declares: name -> (restype, argtypes)
_vp, _sz, _i, _d, _u64 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_double, ctypes.c_uint64
_pp = ctypes.POINTER(ctypes.c_void_p)
SIGNATURES = {
    "b200_handle_create": (_i, [_vp, _pp]),
    "b200_handle_create_mg": (_i, [_vp, _vp, _i, _i, _i, _pp]),
    "b200_nccl_unique_id": (_i, [_vp]),
    "b200_handle_rank": (_i, [_vp]),
    "b200_handle_world_size": (_i, [_vp]),
    "b200_handle_free": (None, [_vp]),
    "b200_last_error": (ctypes.c_char_p, []),
    "b200_graph_create_sg": (_i, [_vp, _vp, _vp, _vp, _sz, _i, _i, _vp, _sz, _i, _i, _pp]),
    "b200_graph_create_mg": (_i, [_vp, _vp, _vp, _vp, _sz, _i, _i, _vp, _sz, _i, _pp]),
    "b200_graph_view": (_i, [_vp, ctypes.POINTER(GraphView)]),
    "b200_graph_partition_view": (_i, [_vp, _i, ctypes.POINTER(GraphView)]),
    "b200_graph_vertex_partition_offsets": (_i, [_vp, ctypes.POINTER(ctypes.c_int64)]),
    "b200_graph_free": (None, [_vp]),
    "b200_graph_out_weight_sums": (_i, [_vp, _vp, _i, _vp]),
    "b200_graph_in_degrees": (_i, [_vp, _vp, _vp]),
    "b200_spmv_pull": (_i, [_vp, _vp, _i, _vp, _d, _d, _vp]),
    "b200_spmv_pull_noalloc": (_i, [_vp, _vp, _i, _vp, _d, _d, _vp]),
    "b200_pagerank": (_i, [_vp, _vp, ctypes.POINTER(PagerankParams), _vp,
                           ctypes.POINTER(CentralityStats)]),
    "b200_katz_centrality": (_i, [_vp, _vp, ctypes.POINTER(KatzParams), _vp,
                                  ctypes.POINTER(CentralityStats)]),
    "b200_eigenvector_centrality": (_i, [_vp, _vp, ctypes.POINTER(EigenvectorParams), _vp,
                                         ctypes.POINTER(CentralityStats)]),
    "b200_hits": (_i, [_vp, _vp, ctypes.POINTER(HitsParams), _vp, _vp,
                       ctypes.POINTER(CentralityStats)]),
    "b200_generate_rmat_edgelist": (_i, [_vp, _i, _sz, _d, _d, _d, _u64, _i, _u64, _i, _vp, _vp]),
    "b200_uniform_fill": (_i, [_vp, _i, _vp, _sz, _u64]),
    "b200_read_matrix_market": (_i, [ctypes.c_char_p, _i, _i, _i, _pp]),
    "b200_read_csv_edgelist": (_i, [ctypes.c_char_p, ctypes.c_char, _i, _i, _i, _i, _pp]),
    "b200_host_edgelist_info": (_i, [_vp, ctypes.POINTER(HostEdgelistInfo)]),
    "b200_host_edgelist_src": (_vp, [_vp]),
    "b200_host_edgelist_dst": (_vp, [_vp]),
    "b200_host_edgelist_weights": (_vp, [_vp]),
    "b200_host_edgelist_free": (None, [_vp]),
    "b200_read_matrix_market_device": (_i, [_vp, ctypes.c_char_p, _i, _i, _pp]),
    "b200_read_csv_edgelist_device": (_i, [_vp, ctypes.c_char_p, ctypes.c_char, _i, _i, _i, _pp]),
    "b200_device_edgelist_info": (_i, [_vp, ctypes.POINTER(HostEdgelistInfo)]),
    "b200_device_edgelist_src": (_vp, [_vp]),
    "b200_device_edgelist_dst": (_vp, [_vp]),
    "b200_device_edgelist_weights": (_vp, [_vp]),
    "b200_device_edgelist_free": (None, [_vp]),
    "b200_build_info": (ctypes.c_char_p, []),
}

_lib = None


class CugraphError(RuntimeError):
    """Raised where pylibcugraph's assert_success raises (pylibcugraph/utils.pyx)."""

    def __init__(self, code: int, message: str):
        super().__init__(f"{ERROR_NAMES.get(code, code)}: {message}")
        self.code = code
        self.message = message


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: the CUDA extension is the product and there is no "
                "fallback. Build it with `make -C cugraph_forked_b200/csrc`.")
        # torch's bundled NCCL must be the one libnccl.so.2 resolves to inside a torch process
        try:
            import torch  # noqa: F401
        except Exception:  # pragma: no cover
            pass
        l = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(code: int) -> None:
    if code != 0:
        raise CugraphError(code, lib().b200_last_error().decode())
