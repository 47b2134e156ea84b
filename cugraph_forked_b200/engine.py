"""Thin object layer over the C ABI: device memory is held in torch CUDA tensors, every compute
call goes through libcugraph_b200.so (cugraph_forked_b200/_lib.py).  Mirrors the C++ API the
reference exposes for this path (cpp/include/cugraph/algorithms.hpp:1147-1159, :1187, :1272-1283;
graph_functions.hpp create_graph_from_edgelist) with internal-order arrays."""
from __future__ import annotations

import ctypes
import weakref
from dataclasses import dataclass

import torch

from . import _lib
from ._lib import FLOAT32, FLOAT64, INT32, INT64, CugraphError, check  # noqa: F401

_DT = {torch.int32: INT32, torch.int64: INT64, torch.float32: FLOAT32, torch.float64: FLOAT64}
_TD = {v: k for k, v in _DT.items()}


def build_info() -> str:
    """the library's version / kernel-generation tag (b200_build_info)"""
    return _lib.lib().b200_build_info().decode()


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _dev(t, dtype=None):
    if t is None:
        return None
    if not isinstance(t, torch.Tensor):
        t = torch.as_tensor(t)
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    if not t.is_cuda:
        t = t.cuda()
    return t.contiguous()


@dataclass
class Stats:
    iterations: int
    converged: bool
    last_diff: float
    elapsed_ms: float
    gpu_launches: int
    spmv_ms: float = 0.0
    spmv_launches: int = 0


class Handle:
    """raft::handle_t stand-in: a stream (+ NCCL communicators when multi-GPU)."""

    def __init__(self, stream: torch.cuda.Stream | None = None):
        self._h = ctypes.c_void_p()
        self.stream = stream
        sp = ctypes.c_void_p(stream.cuda_stream) if stream is not None else None
        check(_lib.lib().b200_handle_create(sp, ctypes.byref(self._h)))

    @classmethod
    def multi_gpu(cls, unique_id: bytes, rank: int, world_size: int, row_comm_size: int,
                  stream: torch.cuda.Stream | None = None) -> "Handle":
        self = cls.__new__(cls)
        self._h = ctypes.c_void_p()
        self.stream = stream
        sp = ctypes.c_void_p(stream.cuda_stream) if stream is not None else None
        buf = ctypes.create_string_buffer(unique_id, 128)
        check(_lib.lib().b200_handle_create_mg(sp, buf, rank, world_size, row_comm_size,
                                               ctypes.byref(self._h)))
        return self

    @staticmethod
    def nccl_unique_id() -> bytes:
        buf = ctypes.create_string_buffer(128)
        check(_lib.lib().b200_nccl_unique_id(buf))
        return buf.raw

    @property
    def rank(self) -> int:
        return _lib.lib().b200_handle_rank(self._h)

    @property
    def world_size(self) -> int:
        return _lib.lib().b200_handle_world_size(self._h)

    def _track(self, graph) -> None:
        # graphs free their device buffers on the handle's stream: they must go first
        if not hasattr(self, "_graphs"):
            self._graphs = weakref.WeakSet()
        self._graphs.add(graph)

    def close(self):
        if self._h:
            for g in list(getattr(self, "_graphs", ())):
                g.close()
            _lib.lib().b200_handle_free(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Graph:
    """graph_t<vertex_t, edge_t, weight_t, store_transposed, multi_gpu> stand-in."""

    def __init__(self, handle: Handle, src, dst, weights=None, *, vertices=None,
                 store_transposed: bool = True, renumber: bool = True, multi_gpu: bool = False):
        self.handle = handle
        src = _dev(src)
        dst = _dev(dst, src.dtype)
        assert src.dtype in (torch.int32, torch.int64) and src.shape == dst.shape
        if weights is not None:
            weights = _dev(weights)
            assert weights.dtype in (torch.float32, torch.float64)
        vertices = _dev(vertices, src.dtype)
        self._g = ctypes.c_void_p()
        wt = _DT[weights.dtype] if weights is not None else FLOAT32
        torch.cuda.current_stream().synchronize()
        if multi_gpu:
            check(_lib.lib().b200_graph_create_mg(
                handle._h, _ptr(src), _ptr(dst), _ptr(weights), src.numel(), _DT[src.dtype], wt,
                _ptr(vertices), 0 if vertices is None else vertices.numel(),
                int(store_transposed), ctypes.byref(self._g)))
        else:
            check(_lib.lib().b200_graph_create_sg(
                handle._h, _ptr(src), _ptr(dst), _ptr(weights), src.numel(), _DT[src.dtype], wt,
                _ptr(vertices), 0 if vertices is None else vertices.numel(),
                int(store_transposed), int(renumber), ctypes.byref(self._g)))
        self.multi_gpu = multi_gpu
        self.vertex_dtype = src.dtype
        handle._track(self)

    # -- introspection (bit-exact parity tests read these) --
    def view(self, partition: int | None = None) -> _lib.GraphView:
        v = _lib.GraphView()
        if partition is None:
            check(_lib.lib().b200_graph_view(self._g, ctypes.byref(v)))
        else:
            check(_lib.lib().b200_graph_partition_view(self._g, partition, ctypes.byref(v)))
        return v

    @staticmethod
    def _copy(ptr, n, dtype):
        """Clone n elements at device pointer `ptr` into a torch tensor."""
        if n == 0:
            return torch.empty(0, dtype=dtype, device="cuda")
        typestr = {torch.int32: "<i4", torch.int64: "<i8", torch.float32: "<f4",
                   torch.float64: "<f8"}[dtype]

        class _View:
            __cuda_array_interface__ = {"shape": (n,), "typestr": typestr,
                                        "data": (int(ptr), False), "version": 2}

        return torch.as_tensor(_View(), device="cuda").clone()

    @property
    def num_vertices(self) -> int:
        return self.view().num_vertices

    @property
    def num_local_vertices(self) -> int:
        v = self.view()
        return v.local_vertex_last - v.local_vertex_first

    @property
    def num_edges(self) -> int:
        return self.view().num_edges

    def segment_offsets(self) -> list[int]:
        v = self.view()
        return [int(v.segment_offsets[i]) for i in range(v.num_segments + 1)] if v.num_segments else []

    def vertex_partition_offsets(self) -> list[int]:
        n = self.handle.world_size if self.multi_gpu else 1
        buf = (ctypes.c_int64 * (n + 1))()
        check(_lib.lib().b200_graph_vertex_partition_offsets(self._g, buf))
        return [int(x) for x in buf]

    def partition_segment_offsets(self, partition: int) -> list[int]:
        v = self.view(partition)
        return [int(v.segment_offsets[i]) for i in range(v.num_segments + 1)]

    def offsets(self, partition=None) -> torch.Tensor:
        v = self.view(partition)
        n = (v.major_last - v.major_first) if partition is not None else v.num_vertices
        return self._copy(v.offsets, n + 1, _TD[v.edge_type])

    def _part_edges(self, partition) -> int:
        # the partition view reports the partition's own edge count
        return self.view(partition).num_edges

    def indices(self, partition=None) -> torch.Tensor:
        """minor OFFSETS (global minor id - view.minor_first), ascending inside every row"""
        v = self.view(partition)
        return self._copy(v.indices, self._part_edges(partition), torch.int32)

    def weights(self, partition=None):
        v = self.view(partition)
        return None if not v.weights else self._copy(v.weights, self._part_edges(partition),
                                                     _TD[v.weight_type])

    def dcs(self, partition: int):
        """(dcs_offsets, dcs_nzd_vertices, major_hypersparse_first) of a multi-GPU edge partition,
        or None when it has no hypersparse segment (compress_hypersparse_offsets)."""
        v = self.view(partition)
        if v.major_hypersparse_first < 0:
            return None
        hs = v.major_hypersparse_first - v.major_first
        return (self._copy(v.dcs_offsets, hs + v.dcs_nzd_count + 1, _TD[v.edge_type]),
                self._copy(v.dcs_nzd_vertices, v.dcs_nzd_count, torch.int32),
                int(v.major_hypersparse_first))

    def renumber_map(self) -> torch.Tensor:
        v = self.view()
        return self._copy(v.renumber_map, v.local_vertex_last - v.local_vertex_first,
                          _TD[v.vertex_type])

    def out_weight_sums(self, dtype=torch.float32) -> torch.Tensor:
        out = torch.empty(self.num_local_vertices, dtype=dtype, device="cuda")
        check(_lib.lib().b200_graph_out_weight_sums(self.handle._h, self._g, _DT[dtype], _ptr(out)))
        return out

    def in_degrees(self) -> torch.Tensor:
        v = self.view()
        out = torch.empty(v.num_vertices, dtype=_TD[v.edge_type], device="cuda")
        check(_lib.lib().b200_graph_in_degrees(self.handle._h, self._g, _ptr(out)))
        return out

    def close(self):
        if self._g:
            _lib.lib().b200_graph_free(self._g)
            self._g = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _stats(s: _lib.CentralityStats) -> Stats:
    return Stats(int(s.iterations), bool(s.converged), float(s.last_diff), float(s.elapsed_ms),
                 int(s.gpu_launches), float(s.spmv_ms), int(s.spmv_launches))


def spmv_pull(graph: Graph, x: torch.Tensor, scale: float = 1.0, init: float = 0.0) -> torch.Tensor:
    """y[dst] = init + scale * sum_{src->dst} x[src] * w  (per_v_transform_reduce_incoming_e)."""
    x = _dev(x)
    y = torch.empty_like(x)
    torch.cuda.current_stream().synchronize()
    check(_lib.lib().b200_spmv_pull(graph.handle._h, graph._g, _DT[x.dtype], _ptr(x), scale, init,
                                    _ptr(y)))
    return y


def pagerank(graph: Graph, *, alpha=0.85, epsilon=1e-6, max_iterations=500, dtype=torch.float32,
             precomputed_out_weight_sums=None, personalization_vertices=None,
             personalization_values=None, initial_guess=None, fixed_iterations=0):
    """cugraph::pagerank on internal-order arrays. Returns (pageranks, Stats)."""
    n = graph.num_local_vertices
    pr = torch.empty(n, dtype=dtype, device="cuda") if initial_guess is None \
        else _dev(initial_guess, dtype).clone()
    ow = _dev(precomputed_out_weight_sums, dtype)
    pv = _dev(personalization_vertices, torch.int32)
    pw = _dev(personalization_values, dtype)
    p = _lib.PagerankParams(alpha, epsilon, max_iterations, _DT[dtype],
                            int(initial_guess is not None), _ptr(ow), _ptr(pv), _ptr(pw),
                            0 if pv is None else pv.numel(), fixed_iterations)
    st = _lib.CentralityStats()
    torch.cuda.current_stream().synchronize()
    check(_lib.lib().b200_pagerank(graph.handle._h, graph._g, ctypes.byref(p), _ptr(pr),
                                   ctypes.byref(st)))
    return pr, _stats(st)


def katz_centrality(graph: Graph, *, alpha=0.01, beta=1.0, betas=None, epsilon=1e-6,
                    max_iterations=500, dtype=torch.float32, initial_guess=None, normalize=False,
                    fixed_iterations=0):
    n = graph.num_local_vertices
    kc = torch.zeros(n, dtype=dtype, device="cuda") if initial_guess is None \
        else _dev(initial_guess, dtype).clone()
    b = _dev(betas, dtype)
    p = _lib.KatzParams(alpha, beta, epsilon, max_iterations, _DT[dtype],
                        int(initial_guess is not None), int(normalize), _ptr(b), fixed_iterations)
    st = _lib.CentralityStats()
    torch.cuda.current_stream().synchronize()
    check(_lib.lib().b200_katz_centrality(graph.handle._h, graph._g, ctypes.byref(p), _ptr(kc),
                                          ctypes.byref(st)))
    return kc, _stats(st)


def eigenvector_centrality(graph: Graph, *, epsilon=1e-6, max_iterations=500, dtype=torch.float32,
                           initial_guess=None, fixed_iterations=0):
    n = graph.num_local_vertices
    c = torch.empty(n, dtype=dtype, device="cuda") if initial_guess is None \
        else _dev(initial_guess, dtype).clone()
    p = _lib.EigenvectorParams(epsilon, max_iterations, _DT[dtype], int(initial_guess is not None),
                               fixed_iterations)
    st = _lib.CentralityStats()
    torch.cuda.current_stream().synchronize()
    check(_lib.lib().b200_eigenvector_centrality(graph.handle._h, graph._g, ctypes.byref(p),
                                                 _ptr(c), ctypes.byref(st)))
    return c, _stats(st)


def hits(graph: Graph, *, epsilon=1e-5, max_iterations=100, dtype=torch.float32,
         initial_hubs_guess=None, normalize=True, fixed_iterations=0):
    """cugraph::hits: returns (hubs, authorities, stats) in internal vertex order;
    stats.iterations is the reference's final_iteration_count, stats.last_diff its hub score
    difference."""
    n = graph.num_local_vertices
    hubs = torch.zeros(n, dtype=dtype, device="cuda") if initial_hubs_guess is None \
        else _dev(initial_hubs_guess, dtype).clone()
    auth = torch.zeros(n, dtype=dtype, device="cuda")
    p = _lib.HitsParams(epsilon, max_iterations, _DT[dtype], int(initial_hubs_guess is not None),
                        int(normalize), 0, fixed_iterations)
    st = _lib.CentralityStats()
    torch.cuda.current_stream().synchronize()
    check(_lib.lib().b200_hits(graph.handle._h, graph._g, ctypes.byref(p), _ptr(hubs), _ptr(auth),
                               ctypes.byref(st)))
    return hubs, auth, _stats(st)


def generate_rmat_edgelist(handle: Handle, scale: int, num_edges: int, a=0.57, b=0.19, c=0.19,
                           seed=0, clip_and_flip=False, first_edge=0, dtype=torch.int32):
    """cugraph::generate_rmat_edgelist: returns (src, dst) device tensors."""
    src = torch.empty(num_edges, dtype=dtype, device="cuda")
    dst = torch.empty(num_edges, dtype=dtype, device="cuda")
    torch.cuda.current_stream().synchronize()
    check(_lib.lib().b200_generate_rmat_edgelist(handle._h, scale, num_edges, a, b, c, seed,
                                                 int(clip_and_flip), first_edge, _DT[dtype],
                                                 _ptr(src), _ptr(dst)))
    return src, dst


def uniform_fill(handle: Handle, n: int, seed=0, dtype=torch.float32) -> torch.Tensor:
    out = torch.empty(n, dtype=dtype, device="cuda")
    torch.cuda.current_stream().synchronize()
    check(_lib.lib().b200_uniform_fill(handle._h, _DT[dtype], _ptr(out), n, seed))
    return out
