"""Edge-list ingest: Matrix-Market and delimited text files -> COO columns.

Host mirrors of the library's readers (include/b200_cugraph.h "Edge-list ingest",
csrc/ingest.cu), which replace
  read_edgelist_from_matrix_market_file  cpp/tests/utilities/matrix_market_file_utilities.cu:266-322
  cudf.read_csv(..., delimiter=' ', dtype=['int32','int32','float32'], header=None)
                                         python/cugraph/cugraph/testing/utils.py:239 read_csv_file,
                                         experimental/datasets/dataset.py:116
The parse runs on the host threads inside the library into page-locked memory; `device="cuda"`
then costs one copy per column.  `device="cpu"` returns the host columns (torch CPU tensors)."""
from __future__ import annotations

import ctypes
import os

import numpy as np
import torch

from . import _lib

_VERTEX = {torch.int32: (_lib.INT32, np.int32), torch.int64: (_lib.INT64, np.int64)}
_WEIGHT = {torch.float32: (_lib.FLOAT32, np.float32), torch.float64: (_lib.FLOAT64, np.float64)}


class EdgeListInfo:
    """What the file said: num_rows / num_cols (MTX size line; CSV: largest id + 1), the number
    of entries in the file and of edges returned (mirrored copies included), is_symmetric
    (graph_properties_t::is_symmetric of read_graph_from_matrix_market_file), is_pattern."""

    def __init__(self, c):
        self.num_rows, self.num_cols = int(c.num_rows), int(c.num_cols)
        self.num_vertices = self.num_rows
        self.num_edges, self.file_entries = int(c.num_edges), int(c.file_entries)
        self.is_symmetric, self.is_pattern = bool(c.is_symmetric), bool(c.is_pattern)
        self.pinned = bool(c.pinned)
        self.parser = "device" if c.parser else "host"

    def __repr__(self):
        return (f"EdgeListInfo(num_vertices={self.num_vertices}, num_edges={self.num_edges}, "
                f"is_symmetric={self.is_symmetric}, is_pattern={self.is_pattern})")


def _column(ptr, n, np_dtype, device):
    if n == 0:
        return torch.empty(0, dtype=torch.from_numpy(np.empty(0, np_dtype)).dtype, device=device)
    nbytes = n * np.dtype(np_dtype).itemsize
    host = np.frombuffer((ctypes.c_char * nbytes).from_address(ptr), dtype=np_dtype)
    if device == "cpu":
        return torch.from_numpy(host.copy())
    out = torch.from_numpy(host).to(device)
    torch.cuda.current_stream().synchronize()  # the host arrays are freed right after
    return out


def _finish_device(el):
    """Columns of a b200_device_edgelist_t: copied into torch-owned tensors, library arrays freed."""
    lib = _lib.lib()
    try:
        info = _lib.HostEdgelistInfo()
        _lib.check(lib.b200_device_edgelist_info(el, ctypes.byref(info)))
        n = int(info.num_edges)

        def column(ptr, dtype):
            out = torch.empty(n, dtype=dtype, device="cuda")
            if n:
                class _Raw:
                    pass
                raw = _Raw()
                raw.__cuda_array_interface__ = {
                    "shape": (n,), "typestr": {torch.int32: "<i4", torch.int64: "<i8",
                                               torch.float32: "<f4", torch.float64: "<f8"}[dtype],
                    "data": (ptr, False), "version": 2}
                out.copy_(torch.as_tensor(raw, device="cuda"))
            return out

        vdt = torch.int32 if info.vertex_type == _lib.INT32 else torch.int64
        src = column(lib.b200_device_edgelist_src(el), vdt)
        dst = column(lib.b200_device_edgelist_dst(el), vdt)
        w = None
        if info.weight_type >= 0:
            wdt = torch.float32 if info.weight_type == _lib.FLOAT32 else torch.float64
            w = column(lib.b200_device_edgelist_weights(el), wdt)
        torch.cuda.current_stream().synchronize()
        return src, dst, w, EdgeListInfo(info)
    finally:
        lib.b200_device_edgelist_free(el)


def _handle_ptr(handle):
    """An engine.Handle (its library stream and device) for the device parser."""
    if handle is None:
        from . import engine
        handle = engine.Handle()
    return handle, handle._h


def _finish(el, device):
    lib = _lib.lib()
    try:
        info = _lib.HostEdgelistInfo()
        _lib.check(lib.b200_host_edgelist_info(el, ctypes.byref(info)))
        n = int(info.num_edges)
        vdt = np.int32 if info.vertex_type == _lib.INT32 else np.int64
        src = _column(lib.b200_host_edgelist_src(el), n, vdt, device)
        dst = _column(lib.b200_host_edgelist_dst(el), n, vdt, device)
        w = None
        if info.weight_type >= 0:
            wdt = np.float32 if info.weight_type == _lib.FLOAT32 else np.float64
            w = _column(lib.b200_host_edgelist_weights(el), n, wdt, device)
        return src, dst, w, EdgeListInfo(info)
    finally:
        lib.b200_host_edgelist_free(el)


def _types(vertex_dtype, weight_dtype):
    if vertex_dtype not in _VERTEX:
        raise TypeError("vertex_dtype must be torch.int32 or torch.int64")
    if weight_dtype is not None and weight_dtype not in _WEIGHT:
        raise TypeError("weight_dtype must be torch.float32, torch.float64 or None")
    return _VERTEX[vertex_dtype][0], (-1 if weight_dtype is None else _WEIGHT[weight_dtype][0])


def read_matrix_market(path, vertex_dtype=torch.int32, weight_dtype=torch.float32, device="cuda",
                       threads=0, parser="host", handle=None):
    """Matrix-Market coordinate file -> (src, dst, weights, info).  Ids become 0-based; a
    symmetric file yields both directions of every off-diagonal entry, the mirrored copy right
    after the entry (mm_to_coo :199-215); pattern files get weight 1.0; weight_dtype=None drops
    the values (the reference's test_weighted=false).  parser="device" parses the text on the GPU
    (csrc/ingest_gpu.cu; same result) and needs device="cuda"."""
    vt, wt = _types(vertex_dtype, weight_dtype)
    el = ctypes.c_void_p()
    if parser == "device":
        if device != "cuda":
            raise ValueError('parser="device" returns CUDA columns: device must be "cuda"')
        handle, hp = _handle_ptr(handle)
        _lib.check(_lib.lib().b200_read_matrix_market_device(hp, os.fsencode(path), vt, wt,
                                                             ctypes.byref(el)))
        return _finish_device(el)
    if parser != "host":
        raise ValueError('parser must be "host" or "device"')
    _lib.check(_lib.lib().b200_read_matrix_market(os.fsencode(path), vt, wt, int(threads),
                                                  ctypes.byref(el)))
    return _finish(el, device)


def read_csv(path, delimiter=" ", vertex_dtype=torch.int32, weight_dtype=torch.float32,
             skip_rows=0, device="cuda", threads=0, parser="host", handle=None):
    """`src<delim>dst[<delim>weight]` lines -> (src, dst, weights, info); ids as written.
    weight_dtype=None reads two columns and ignores anything after them.  parser="device": as in
    read_matrix_market."""
    if not isinstance(delimiter, str) or len(delimiter) != 1:
        raise ValueError("delimiter must be one character")
    vt, wt = _types(vertex_dtype, weight_dtype)
    el = ctypes.c_void_p()
    if parser == "device":
        if device != "cuda":
            raise ValueError('parser="device" returns CUDA columns: device must be "cuda"')
        handle, hp = _handle_ptr(handle)
        _lib.check(_lib.lib().b200_read_csv_edgelist_device(hp, os.fsencode(path),
                                                            delimiter.encode(), int(skip_rows),
                                                            vt, wt, ctypes.byref(el)))
        return _finish_device(el)
    if parser != "host":
        raise ValueError('parser must be "host" or "device"')
    _lib.check(_lib.lib().b200_read_csv_edgelist(os.fsencode(path), delimiter.encode(),
                                                 int(skip_rows), vt, wt, int(threads),
                                                 ctypes.byref(el)))
    return _finish(el, device)
