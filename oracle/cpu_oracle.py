"""TEST INFRASTRUCTURE ONLY — ctypes loader for the CPU oracles.

  libcentrality_oracle.so        our restatement   (oracle/centrality_oracle.cpp)
  _ref/libcugraph_ref_oracle.so  the reference's own host functions (oracle/ref_shim.cpp +
                                 oracle/build_ref.sh), present when built in the container.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PORT = os.path.join(_HERE, "libcentrality_oracle.so")
_REF = os.path.join(_HERE, "_ref", "libcugraph_ref_oracle.so")
_libs: dict[str, ctypes.CDLL] = {}


def build(force: bool = False) -> None:
    if force or not os.path.exists(_PORT) or any(
        os.path.getmtime(_PORT) < os.path.getmtime(os.path.join(_HERE, f))
        for f in ("centrality_oracle.cpp", "inputs_oracle.cpp")
    ):
        subprocess.check_call(["bash", os.path.join(_HERE, "build_ref.sh")])


def port_lib() -> ctypes.CDLL:
    if "port" not in _libs:
        build()
        _libs["port"] = ctypes.CDLL(_PORT)
    return _libs["port"]


def ref_available() -> bool:
    return os.path.exists(_REF)


def ref_lib() -> ctypes.CDLL:
    if "ref" not in _libs:
        _libs["ref"] = ctypes.CDLL(_REF)
    return _libs["ref"]


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _sfx(dtype):
    return {np.dtype(np.float32): "f32", np.dtype(np.float64): "f64"}[np.dtype(dtype)]


def _prep(offsets, indices, weights, dtype):
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    indices = np.ascontiguousarray(indices, dtype=np.int32)
    weights = None if weights is None else np.ascontiguousarray(weights, dtype=dtype)
    return offsets, indices, weights


def pagerank(offsets, indices, weights=None, *, dtype=np.float32, alpha=0.85, epsilon=1e-6,
             max_iterations=500, personalization=None, initial_guess=None, fixed_iterations=0,
             impl="port"):
    """Returns (scores, iterations). impl="ref" runs the reference's own pagerank_reference."""
    offsets, indices, weights = _prep(offsets, indices, weights, dtype)
    nv = len(offsets) - 1
    pr = np.zeros(nv, dtype=dtype) if initial_guess is None else np.array(initial_guess, dtype=dtype)
    pv = pw = None
    if personalization is not None:
        pv = np.ascontiguousarray(personalization[0], dtype=np.int32)
        pw = np.ascontiguousarray(personalization[1], dtype=dtype)
    npers = 0 if pv is None else len(pv)
    if impl == "ref":
        fn = getattr(ref_lib(), "ref_pagerank_" + _sfx(dtype))
        fn.restype = ctypes.c_long
        if fixed_iterations > 0:
            # pagerank_reference has no iteration pin of its own: with epsilon = 0 its stop test
            # (diff_sum < epsilon, pagerank_test.cpp:124) never holds, so it runs exactly
            # max_iterations iterations, leaves their result in `pr`, and then trips its
            # ASSERT_TRUE(iter < max_iterations), which the shim reports as -1.
            rc = fn(_p(offsets), _p(indices), _p(weights), _p(pv), _p(pw), ctypes.c_int64(npers),
                    _p(pr), ctypes.c_int64(nv), ctypes.c_double(alpha), ctypes.c_double(0.0),
                    ctypes.c_long(fixed_iterations), ctypes.c_int(initial_guess is not None))
            if rc != -1:
                raise RuntimeError("pagerank_reference did not run to the pinned iteration count")
            return pr, int(fixed_iterations)
        rc = fn(_p(offsets), _p(indices), _p(weights), _p(pv), _p(pw), ctypes.c_int64(npers), _p(pr),
                ctypes.c_int64(nv), ctypes.c_double(alpha), ctypes.c_double(epsilon),
                ctypes.c_long(max_iterations), ctypes.c_int(initial_guess is not None))
    else:
        # impl="ext": the port with its sums in extended precision (fp64 values only)
        name = "oracle_pagerank_f64_ext" if impl == "ext" else "oracle_pagerank_" + _sfx(dtype)
        assert impl != "ext" or np.dtype(dtype) == np.float64
        fn = getattr(port_lib(), name)
        fn.restype = ctypes.c_long
        rc = fn(_p(offsets), _p(indices), _p(weights), _p(pv), _p(pw), ctypes.c_int64(npers), _p(pr),
                ctypes.c_int64(nv), ctypes.c_double(alpha), ctypes.c_double(epsilon),
                ctypes.c_long(max_iterations), ctypes.c_int(initial_guess is not None),
                ctypes.c_long(fixed_iterations))
    if rc < 0:
        raise RuntimeError("PageRank failed to converge." if rc == -1 else "invalid input")
    return pr, int(rc)


def katz(offsets, indices, weights=None, *, dtype=np.float32, alpha=0.01, beta=1.0, betas=None,
         epsilon=1e-6, max_iterations=500, initial_guess=None, normalize=False,
         fixed_iterations=0, impl="port"):
    offsets, indices, weights = _prep(offsets, indices, weights, dtype)
    nv = len(offsets) - 1
    kc = np.zeros(nv, dtype=dtype) if initial_guess is None else np.array(initial_guess, dtype=dtype)
    b = None if betas is None else np.ascontiguousarray(betas, dtype=dtype)
    if impl == "ref":
        fn = getattr(ref_lib(), "ref_katz_" + _sfx(dtype))
        fn.restype = ctypes.c_long
        rc = fn(_p(offsets), _p(indices), _p(weights), _p(b), _p(kc), ctypes.c_int64(nv),
                ctypes.c_double(alpha), ctypes.c_double(beta), ctypes.c_double(epsilon),
                ctypes.c_long(max_iterations), ctypes.c_int(initial_guess is not None),
                ctypes.c_int(normalize))
    else:
        fn = getattr(port_lib(), "oracle_katz_" + _sfx(dtype))
        fn.restype = ctypes.c_long
        rc = fn(_p(offsets), _p(indices), _p(weights), _p(b), _p(kc), ctypes.c_int64(nv),
                ctypes.c_double(alpha), ctypes.c_double(beta), ctypes.c_double(epsilon),
                ctypes.c_long(max_iterations), ctypes.c_int(initial_guess is not None),
                ctypes.c_int(normalize), ctypes.c_long(fixed_iterations))
    if rc < 0:
        raise RuntimeError("Katz Centrality failed to converge.")
    return kc, int(rc)


def eigenvector(src, dst, weights=None, *, num_vertices, dtype=np.float32, epsilon=1e-6,
                max_iterations=500, fixed_iterations=0, impl="port"):
    """NB the reference oracle accumulates into tmp[src[e]] from old[dst[e]] (the test feeds it
    the edge list in the orientation that makes this the pull direction)."""
    src = np.ascontiguousarray(src, dtype=np.int32)
    dst = np.ascontiguousarray(dst, dtype=np.int32)
    w = None if weights is None else np.ascontiguousarray(weights, dtype=dtype)
    out = np.zeros(num_vertices, dtype=dtype)
    if impl == "ref":
        fn = getattr(ref_lib(), "ref_eigenvector_" + _sfx(dtype))
        fn.restype = ctypes.c_long
        rc = fn(_p(src), _p(dst), _p(w), ctypes.c_int64(len(src)), _p(out),
                ctypes.c_int64(num_vertices), ctypes.c_double(epsilon), ctypes.c_long(max_iterations))
    else:
        fn = getattr(port_lib(), "oracle_eigenvector_" + _sfx(dtype))
        fn.restype = ctypes.c_long
        rc = fn(_p(src), _p(dst), _p(w), ctypes.c_int64(len(src)), _p(out),
                ctypes.c_int64(num_vertices), ctypes.c_double(epsilon), ctypes.c_long(max_iterations),
                ctypes.c_long(fixed_iterations))
    if rc < 0:
        raise RuntimeError("Eigenvector Centrality failed to converge.")
    return out, int(rc)


def hits(offsets, indices, *, dtype=np.float32, max_iterations=500, tolerance=1e-5,
         starting_hubs=None, normalize=True, fixed_iterations=0, impl="port"):
    """HITS on the CSR (rows = sources) of a graph.  Returns (hubs, authorities, error, iterations).
    impl="ref" runs the reference's own hits_reference (always sum-normalised)."""
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    indices = np.ascontiguousarray(indices, dtype=np.int32)
    nv = len(offsets) - 1
    hubs = np.zeros(nv, dtype=dtype)
    auth = np.zeros(nv, dtype=dtype)
    err = ctypes.c_double(0.0)
    sh = None if starting_hubs is None else np.ascontiguousarray(starting_hubs, dtype=dtype)
    if impl == "ref":
        assert normalize and not fixed_iterations
        fn = getattr(ref_lib(), "ref_hits_" + _sfx(dtype))
        fn.restype = ctypes.c_long
        rc = fn(_p(offsets), _p(indices), ctypes.c_int64(nv), ctypes.c_int64(len(indices)),
                ctypes.c_long(max_iterations), _p(sh), ctypes.c_double(tolerance), _p(hubs),
                _p(auth), ctypes.byref(err))
    else:
        fn = getattr(port_lib(), "oracle_hits_" + _sfx(dtype))
        fn.restype = ctypes.c_long
        rc = fn(_p(offsets), _p(indices), ctypes.c_int64(nv), ctypes.c_long(max_iterations), _p(sh),
                ctypes.c_int(normalize), ctypes.c_double(tolerance), _p(hubs), _p(auth),
                ctypes.byref(err), ctypes.c_long(fixed_iterations))
    if rc < 0:
        raise RuntimeError("HITS: invalid input")
    return hubs, auth, float(err.value), int(rc)


def csr_from_coo(src, dst, num_vertices):
    """Un-renumbered CSR (rows = sources) in input order, the layout hits_reference walks."""
    order = np.argsort(src, kind="stable")
    offsets = np.zeros(num_vertices + 1, dtype=np.int64)
    np.cumsum(np.bincount(src, minlength=num_vertices), out=offsets[1:])
    return offsets, np.asarray(dst)[order].astype(np.int32)


def csc_from_coo(src, dst, weights, num_vertices):
    """Un-renumbered CSC (rows = destinations) in input order, as the reference tests build the
    host graph for pagerank_reference (pagerank_test.cpp:264-300)."""
    order = np.argsort(dst, kind="stable")
    offsets = np.zeros(num_vertices + 1, dtype=np.int64)
    np.cumsum(np.bincount(dst, minlength=num_vertices), out=offsets[1:])
    return offsets, np.asarray(src)[order].astype(np.int32), None if weights is None else np.asarray(weights)[order]


def rmat_edgelist_fast(scale, num_edges, a=0.57, b=0.19, c=0.19, seed=0, first_edge=0, threads=0):
    """graph_oracle.rmat_edgelist (same stream, same bit loop) in C++ on `threads` host threads."""
    lib = port_lib()
    src = np.empty(num_edges, np.int32)
    dst = np.empty(num_edges, np.int32)
    lib.oracle_rmat_edgelist_i32.restype = None
    lib.oracle_rmat_edgelist_i32(ctypes.c_int(scale), ctypes.c_int64(num_edges), ctypes.c_double(a),
                                 ctypes.c_double(b), ctypes.c_double(c), ctypes.c_uint64(seed),
                                 ctypes.c_uint64(first_edge), _p(src), _p(dst),
                                 ctypes.c_int(threads or (os.cpu_count() or 1)))
    return src, dst


def csc_from_coo_fast(src, dst, num_vertices, threads=0):
    """csc_from_coo for unweighted int32 edge lists: stable counting sort by destination in C++."""
    lib = port_lib()
    src = np.ascontiguousarray(src, np.int32)
    dst = np.ascontiguousarray(dst, np.int32)
    offsets = np.empty(num_vertices + 1, np.int64)
    indices = np.empty(len(src), np.int32)
    threads = min(threads or (os.cpu_count() or 1), 16)
    lib.oracle_csc_from_coo_i32.restype = None
    lib.oracle_csc_from_coo_i32(_p(src), _p(dst), ctypes.c_int64(len(src)),
                                ctypes.c_int64(num_vertices), _p(offsets), _p(indices),
                                ctypes.c_int(threads))
    return offsets, indices


def read_mtx_ref(path, dtype=np.float32):
    """The reference's own reader (mm_properties + mm_to_coo compiled in oracle/_ref), called as
    read_edgelist_from_matrix_market_file does.  Returns (src, dst, weights, num_vertices,
    is_symmetric, is_pattern)."""
    lib = ref_lib()
    m, k, nnz = ctypes.c_int64(), ctypes.c_int64(), ctypes.c_int64()
    sym, pat = ctypes.c_int(), ctypes.c_int()
    lib.ref_mtx_properties.restype = ctypes.c_long
    rc = lib.ref_mtx_properties(os.fsencode(path), ctypes.byref(m), ctypes.byref(k),
                                ctypes.byref(nnz), ctypes.byref(sym), ctypes.byref(pat))
    if rc != 0:
        raise RuntimeError({-1: "fopen failure", -2: "could not read Matrix Market file properties.",
                            -3: "invalid Matrix Market file properties."}[rc])
    rows = np.zeros(nnz.value, np.int32)
    cols = np.zeros(nnz.value, np.int32)
    vals = np.zeros(nnz.value, dtype)
    fn = getattr(lib, "ref_mtx_to_coo_" + _sfx(dtype))
    fn.restype = ctypes.c_long
    if fn(os.fsencode(path), ctypes.c_int64(nnz.value), _p(rows), _p(cols), _p(vals)) != 0:
        raise RuntimeError("could not read matrix data")
    return rows, cols, vals, m.value, bool(sym.value), bool(pat.value)


def graph_ref(src, dst, weights, num_vertices, store_transposed=True):
    """The reference's own host graph build (graph_reference, cpp/tests/structure/graph_test.cpp:
    38-69, compiled in oracle/_ref): (offsets, indices, weights) with the neighbours of a row in
    input order."""
    lib = ref_lib()
    src = np.ascontiguousarray(src, np.int32)
    dst = np.ascontiguousarray(dst, np.int32)
    w = None if weights is None else np.ascontiguousarray(weights, np.float64)
    ne = len(src)
    offsets = np.zeros(num_vertices + 1, np.int64)
    indices = np.zeros(ne, np.int32)
    out_w = np.zeros(ne if w is not None else 0, np.float64)
    lib.ref_graph_i32_f64.restype = ctypes.c_long
    lib.ref_graph_i32_f64(ctypes.c_int(int(store_transposed)), _p(src), _p(dst), _p(w),
                          ctypes.c_int32(num_vertices), ctypes.c_int64(ne), _p(offsets), _p(indices),
                          _p(out_w) if w is not None else None)
    return offsets, indices, (out_w if w is not None else None)


def degree_ref(offsets, indices, num_vertices, major):
    """degree_reference (cpp/tests/structure/degree_test.cpp:37-57)."""
    lib = ref_lib()
    offsets = np.ascontiguousarray(offsets, np.int64)
    indices = np.ascontiguousarray(indices, np.int32)
    out = np.zeros(num_vertices, np.int64)
    lib.ref_degree_i32(_p(offsets), _p(indices), _p(out), ctypes.c_int32(num_vertices),
                       ctypes.c_int(int(major)))
    return out


def weight_sum_ref(offsets, indices, weights, num_vertices, major):
    """weight_sum_reference (cpp/tests/structure/weight_sum_test.cpp:37-59)."""
    lib = ref_lib()
    offsets = np.ascontiguousarray(offsets, np.int64)
    indices = np.ascontiguousarray(indices, np.int32)
    weights = np.ascontiguousarray(weights, np.float64)
    out = np.zeros(num_vertices, np.float64)
    lib.ref_weight_sum_i32_f64(_p(offsets), _p(indices), _p(weights), _p(out),
                               ctypes.c_int32(num_vertices), ctypes.c_int(int(major)))
    return out
