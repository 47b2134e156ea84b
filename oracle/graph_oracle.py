"""TEST INFRASTRUCTURE ONLY — numpy restatement of the reference's integer graph-build path.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module;
nothing under cugraph_forked_b200/ does.  Every function cites the reference file:line it follows
(paths relative to /root/reference).  All outputs are integer artefacts that the CUDA build must
reproduce bit-exactly.

Parity status of third-party arithmetic that is NOT under /root/reference:
  * cuco::detail::MurmurHash3_32 (cuCollections, pinned only as rapids-cmake branch-22.10):
    restated as the published MurmurHash3_x86_32 with seed 0 over the key's 4/8 bytes —
    "parity unpinned" (the reference holds no golden hash values; tests/test_graph_oracle.py
    pins the restatement against the public SMHasher known answers instead).
  * raft::random::uniform: the R-MAT random stream is "parity unpinned"; `rmat_edgelist` uses
    this repository's own counter-based generator (splitmix64 finaliser) and only the bit loop
    follows the reference.
"""
from __future__ import annotations

import numpy as np

# cpp/include/cugraph/graph_view.hpp:255-263
LOW_DEGREE_THRESHOLD = 32
MID_DEGREE_THRESHOLD = 1024
HYPERSPARSE_THRESHOLD_RATIO = 0.5


# ----------------------------------------------------------------------------------------------
# hashing / GPU assignment  (cpp/src/detail/graph_utils.cuh:45-138)
# ----------------------------------------------------------------------------------------------
def murmur3_32(keys: np.ndarray) -> np.ndarray:
    """MurmurHash3_x86_32(seed=0) of each key's raw little-endian bytes (4 for int32, 8 for int64)."""
    keys = np.ascontiguousarray(keys)
    assert keys.dtype in (np.int32, np.int64, np.uint32, np.uint64)
    nbytes = keys.dtype.itemsize
    blocks = keys.view(np.uint32).reshape(-1, nbytes // 4)
    c1, c2 = np.uint32(0xCC9E2D51), np.uint32(0x1B873593)
    h = np.zeros(blocks.shape[0], dtype=np.uint32)
    with np.errstate(over="ignore"):
        for b in range(blocks.shape[1]):
            k = blocks[:, b].copy()
            k *= c1
            k = (k << np.uint32(15)) | (k >> np.uint32(17))
            k *= c2
            h ^= k
            h = (h << np.uint32(13)) | (h >> np.uint32(19))
            h = h * np.uint32(5) + np.uint32(0xE6546B64)
        h ^= np.uint32(nbytes)
        h ^= h >> np.uint32(16)
        h *= np.uint32(0x85EBCA6B)
        h ^= h >> np.uint32(13)
        h *= np.uint32(0xC2B2AE35)
        h ^= h >> np.uint32(16)
    return h


def gpu_id_from_ext_vertex(v: np.ndarray, comm_size: int) -> np.ndarray:
    """compute_gpu_id_from_ext_vertex_t, graph_utils.cuh:45-53."""
    return (murmur3_32(v) % np.uint32(comm_size)).astype(np.int32)


def gpu_id_from_ext_edge_endpoints(major, minor, comm_size, row_comm_size, col_comm_size):
    """compute_gpu_id_from_ext_edge_endpoints_t, graph_utils.cuh:69-82."""
    mj = gpu_id_from_ext_vertex(major, comm_size)
    mn = gpu_id_from_ext_vertex(minor, comm_size)
    return (mn // row_comm_size) * row_comm_size + (mj % row_comm_size)


def partition_id_from_ext_edge_endpoints(major, minor, comm_size, row_comm_size, col_comm_size):
    """compute_partition_id_from_ext_edge_endpoints_t, graph_utils.cuh:124-138."""
    mj = gpu_id_from_ext_vertex(major, comm_size)
    mn = gpu_id_from_ext_vertex(minor, comm_size)
    return mj * col_comm_size + mn // row_comm_size


def grid_2d(comm_size: int) -> tuple[int, int]:
    """(row_comm_size, col_comm_size): row_comm_size = largest divisor <= sqrt(P).
    cpp/tests/utilities/mg_utilities.cpp:58-61; python dask/comms/comms.py:40-44."""
    r = int(np.sqrt(comm_size))
    while comm_size % r != 0:
        r -= 1
    return r, comm_size // r


# ----------------------------------------------------------------------------------------------
# renumbering + degree segments  (cpp/src/structure/renumber_edgelist_impl.cuh:96-452)
# ----------------------------------------------------------------------------------------------
def segment_thresholds(col_comm_size: int | None) -> list[int]:
    """renumber_edgelist_impl.cuh:394-424. None = single GPU."""
    if col_comm_size is None:
        return [MID_DEGREE_THRESHOLD, LOW_DEGREE_THRESHOLD, 1]
    mid = MID_DEGREE_THRESHOLD * col_comm_size
    low = LOW_DEGREE_THRESHOLD * col_comm_size
    hyp = int(col_comm_size * HYPERSPARSE_THRESHOLD_RATIO)
    if hyp > 0:
        return [mid, low, hyp, min(hyp, 1)]
    return [mid, low, 1]


def sort_by_degree_and_segment(sorted_vertices, degrees, thresholds):
    """renumber_edgelist_impl.cuh:384-449: stable descending sort by degree (ties keep ascending
    external id) then upper_bound(greater) against the thresholds."""
    order = np.argsort(-degrees.astype(np.int64), kind="stable")
    labels = sorted_vertices[order]
    d = degrees[order]
    seg = [0] + [int(np.count_nonzero(d >= t)) for t in thresholds] + [int(len(d))]
    return labels, d, seg


def renumber_sg(majors, minors, vertices=None):
    """Single-GPU compute_renumber_map + renumber_edgelist (renumber_edgelist_impl.cuh:96-452,
    :923-994). Returns (renumber_map_labels, segment_offsets, new_majors, new_minors)."""
    if vertices is None:
        sorted_vertices = np.unique(np.concatenate([majors, minors]))
    else:
        sorted_vertices = np.sort(np.asarray(vertices))
    pos = np.searchsorted(sorted_vertices, majors)
    degrees = np.bincount(pos, minlength=len(sorted_vertices)).astype(np.int64)
    labels, _, seg = sort_by_degree_and_segment(sorted_vertices, degrees, segment_thresholds(None))
    new_of_sorted = np.empty(len(labels), dtype=majors.dtype)
    new_of_sorted[np.searchsorted(sorted_vertices, labels)] = np.arange(len(labels), dtype=majors.dtype)
    new_majors = new_of_sorted[pos]
    new_minors = new_of_sorted[np.searchsorted(sorted_vertices, minors)]
    return labels.astype(majors.dtype), seg, new_majors, new_minors


# ----------------------------------------------------------------------------------------------
# COO -> CSC with sorted neighbours  (cpp/src/structure/detail/structure_utils.cuh:173-490)
# ----------------------------------------------------------------------------------------------
def compress_sorted(majors, minors, weights, major_first, major_last):
    """compress_edgelist + sort_adjacency_list for one rectangle. `indices` ascending inside each
    row (the bit-pinned part); weights follow a stable order (the reference's order among equal
    (major, minor) pairs is unspecified — cub segmented sort is not stable)."""
    n = int(major_last - major_first)
    order = np.lexsort((minors, majors))  # stable: by major, then minor, then input order
    counts = np.bincount(majors - major_first, minlength=n) if len(majors) else np.zeros(n, np.int64)
    offsets = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(counts, out=offsets[1:])
    indices = minors[order]
    w = None if weights is None else weights[order]
    return offsets, indices, w


def out_weight_sums(indices, weights, num_minors, dtype):
    """graph_view_t::compute_out_weight_sums (graph_view_impl.cuh:238-265): per-source sum of
    edge weights (out-degree when unweighted). Float sums are order-dependent; compared with
    tolerance, exact when unweighted."""
    if weights is None:
        return np.bincount(indices, minlength=num_minors).astype(dtype)
    return np.bincount(indices, weights=weights.astype(np.float64), minlength=num_minors).astype(dtype)


# ----------------------------------------------------------------------------------------------
# multi-GPU build on VIRTUAL ranks  (create_graph_from_edgelist_impl.cuh:207-555,
# renumber_edgelist_impl.cuh:665-921, graph_view.hpp:81-242)
# ----------------------------------------------------------------------------------------------
def rank_coords(rank: int, row_comm_size: int) -> tuple[int, int]:
    """(row_comm_rank, col_comm_rank) of a world rank: ranks of one row communicator are
    consecutive (partition_manager.hpp:92-100: row colour = rank / R, col colour = rank % R)."""
    return rank % row_comm_size, rank // row_comm_size


def build_mg(src, dst, weights, comm_size, store_transposed=True, vertices=None):
    """What every rank of a P = R x C grid holds after cugraph_mg_graph_create, computed on one
    host.  Returns a dict:
      R, C, num_vertices, vertex_partition_offsets [P+1]
      labels[r]            renumber_map_labels of rank r (external ids in new-id order)
      segment_offsets[r]   vertex-partition segment offsets of rank r (thresholds scaled by C,
                           renumber_edgelist_impl.cuh:394-449)
      edge_rank            owner rank of every input edge (graph_utils.cuh:69-82)
      new_major, new_minor renumbered endpoints of every input edge
      parts[g][i]          edge partition i of rank g: offsets (full, not DCSC-compressed),
                           indices (GLOBAL minor ids, ascending per row), weights, major_first,
                           major_last, minor_first, minor_last, segment_offsets
    Edge partition i of rank g holds the edges whose major is owned by rank R*i + row_rank(g)
    (graph_view.hpp:181-189) and whose minor is owned by a rank of column col_rank(g)
    (:204-212)."""
    P = int(comm_size)
    R, C = grid_2d(P)
    majors, minors = (dst, src) if store_transposed else (src, dst)
    majors = np.asarray(majors)
    minors = np.asarray(minors)
    if vertices is None:
        all_v = np.unique(np.concatenate([majors, minors]))
    else:
        all_v = np.unique(np.asarray(vertices, majors.dtype))
    owner = gpu_id_from_ext_vertex(all_v, P) if len(all_v) else np.zeros(0, np.int32)
    pos_major = np.searchsorted(all_v, majors)
    pos_minor = np.searchsorted(all_v, minors)
    degrees = np.bincount(pos_major, minlength=len(all_v)).astype(np.int64)
    thresholds = segment_thresholds(C)
    labels, segs, counts = [], [], []
    new_of = np.zeros(len(all_v), dtype=np.int64)
    first = 0
    for r in range(P):
        sel = np.nonzero(owner == r)[0]          # ascending external id
        lab, _, seg = sort_by_degree_and_segment(all_v[sel], degrees[sel], thresholds)
        order = np.argsort(-degrees[sel], kind="stable")
        new_of[sel[order]] = first + np.arange(len(sel))
        labels.append(lab.astype(majors.dtype))
        segs.append(seg)
        counts.append(len(sel))
        first += len(sel)
    vpo = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    new_major = new_of[pos_major]
    new_minor = new_of[pos_minor]
    edge_rank = (gpu_id_from_ext_edge_endpoints(majors, minors, P, R, C)
                 if len(majors) else np.zeros(0, np.int32))
    major_owner = owner[pos_major] if len(majors) else np.zeros(0, np.int32)
    parts = []
    for g in range(P):
        row_rank, col_rank = rank_coords(g, R)
        mine = edge_rank == g
        minor_first, minor_last = int(vpo[col_rank * R]), int(vpo[(col_rank + 1) * R])
        plist = []
        for i in range(C):
            vp = R * i + row_rank
            sel = np.nonzero(mine & (major_owner == vp))[0]
            mf, ml = int(vpo[vp]), int(vpo[vp + 1])
            w = None if weights is None else np.asarray(weights)[sel]
            offsets, indices, ww = compress_sorted(new_major[sel], new_minor[sel], w, mf, ml)
            assert len(indices) == 0 or (indices.min() >= minor_first and indices.max() < minor_last)
            plist.append(dict(offsets=offsets, indices=indices, weights=ww, major_first=mf,
                              major_last=ml, minor_first=minor_first, minor_last=minor_last,
                              segment_offsets=segs[vp]))
        parts.append(plist)
    return dict(R=R, C=C, num_vertices=int(vpo[-1]), vertex_partition_offsets=vpo, labels=labels,
                segment_offsets=segs, edge_rank=edge_rank, new_major=new_major,
                new_minor=new_minor, parts=parts)


def compress_hypersparse(offsets, major_first, segment_offsets):
    """compress_hypersparse_offsets, structure_utils.cuh:109-159: rows at or after the start of the
    hypersparse segment keep only their non-empty entries.  Returns (dcs_offsets,
    dcs_nzd_vertices) or None when the vertex partition has no hypersparse segment
    (segment_offsets has 5 entries + end only when col_comm_size >= 2)."""
    if len(segment_offsets) != 6:
        return None
    hs_first = segment_offsets[3]
    deg = np.diff(offsets)
    nz = np.nonzero(deg[hs_first:] > 0)[0] + hs_first
    dcs_offsets = np.concatenate([offsets[:hs_first + 1], offsets[nz + 1]])
    return dcs_offsets, (nz + major_first)


# ----------------------------------------------------------------------------------------------
# R-MAT  (cpp/src/generators/generate_rmat_edgelist.cu:35-111)
# ----------------------------------------------------------------------------------------------
def splitmix64(x: np.ndarray) -> np.ndarray:
    """This repository's counter-based generator (NOT raft's): splitmix64 finaliser of (x)."""
    with np.errstate(over="ignore"):
        z = x.astype(np.uint64) + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def uniform01(counter: np.ndarray, seed: int) -> np.ndarray:
    """float32 in [0,1): top 24 bits of splitmix64(seed + counter) * 2^-24."""
    with np.errstate(over="ignore"):
        z = splitmix64(counter.astype(np.uint64) + np.uint64(seed))
    return ((z >> np.uint64(40)).astype(np.float32)) * np.float32(2.0 ** -24)


def rmat_edgelist(scale, num_edges, a=0.57, b=0.19, c=0.19, seed=0, clip_and_flip=False,
                  first_edge=0, dtype=np.int32):
    """Bit loop of generate_rmat_edgelist.cu:75-107. Random number (edge i, bit, k) has the global
    counter (first_edge + i) * 2 * scale + 2 * bit + k — the reference's rands[] indexing with its
    per-chunk seed advance (`seed += n * 2 * scale`, :69) folded into one global counter."""
    a_plus_b = a + b
    a_norm = a / (a + b) if (a + b) > 0.0 else 0.0
    c_norm = c / (1.0 - (a + b)) if (1.0 - (a + b)) > 0.0 else 0.0
    i = np.arange(first_edge, first_edge + num_edges, dtype=np.uint64)
    src = np.zeros(num_edges, dtype=np.int64)
    dst = np.zeros(num_edges, dtype=np.int64)
    for bit in range(scale - 1, -1, -1):
        base = i * np.uint64(2 * scale) + np.uint64(2 * bit)
        r0 = uniform01(base, seed).astype(np.float64)
        r1 = uniform01(base + np.uint64(1), seed).astype(np.float64)
        sb = r0 > a_plus_b
        db = r1 > np.where(sb, c_norm, a_norm)
        if clip_and_flip:
            flip = (src == dst) & (~sb) & db
            sb = np.where(flip, ~sb, sb)
            db = np.where(flip, ~db, db)
        src += sb.astype(np.int64) << bit
        dst += db.astype(np.int64) << bit
    return src.astype(dtype), dst.astype(dtype)


# ----------------------------------------------------------------------------------------------
# Matrix Market  (cpp/tests/utilities/matrix_market_file_utilities.cu:134-215)
# ----------------------------------------------------------------------------------------------
def read_mtx(path: str, dtype=np.float32):
    """MTX coordinate file -> (src, dst, weights, num_vertices, is_symmetric); 1-based -> 0-based,
    the symmetric complement of every off-diagonal entry is appended right after the entry
    (mm_to_coo, :166-215). Pattern files get weight 1.0."""
    with open(path) as f:
        banner = f.readline().lower().split()
        assert banner[0] == "%%matrixmarket" and banner[2] == "coordinate"
        pattern = banner[3] == "pattern"
        general = banner[4] == "general"
        skew = banner[4] == "skew-symmetric"
        line = f.readline()
        while line.startswith("%"):
            line = f.readline()
        m, n, nnz = (int(t) for t in line.split())
        rows, cols, vals = [], [], []
        for _ in range(nnz):
            t = f.readline().split()
            r, c = int(t[0]) - 1, int(t[1]) - 1
            v = 1.0 if pattern else float(t[2])
            rows.append(r); cols.append(c); vals.append(v)
            if not general and r != c:
                rows.append(c); cols.append(r); vals.append(-v if skew else v)
    return (np.asarray(rows, np.int32), np.asarray(cols, np.int32), np.asarray(vals, dtype),
            max(m, n), not general)
