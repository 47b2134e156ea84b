"""Shared helpers of the parity tests: inputs for the CPU oracle and comparisons."""
import os

import numpy as np

from oracle import cpu_oracle, graph_oracle

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_mtx(name, dtype=np.float32):
    return graph_oracle.read_mtx(os.path.join(GOLDEN, name), dtype)


def rmat(scale, edge_factor=16, seed=0, dtype=np.int32):
    ne = edge_factor << scale
    return graph_oracle.rmat_edgelist(scale, ne, seed=seed, dtype=dtype)


def to_external(values_internal, renumber_map, num_vertices):
    """scores in internal order -> array indexed by external id (ids must be < num_vertices)."""
    out = np.full(num_vertices, np.nan, dtype=values_internal.dtype)
    out[renumber_map] = values_internal
    return out


def rel_err(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), np.finfo(np.float64).tiny)


def top_k(scores, k=100):
    """ids of the k largest scores, ties broken by id (deterministic)."""
    scores = np.asarray(scores)
    order = np.lexsort((np.arange(len(scores)), -scores.astype(np.float64)))
    return order[:k]


def oracle_pagerank(src, dst, w, nv, **kw):
    offsets, indices, weights = cpu_oracle.csc_from_coo(src, dst, w, nv)
    return cpu_oracle.pagerank(offsets, indices, weights, **kw)


def oracle_katz(src, dst, w, nv, **kw):
    offsets, indices, weights = cpu_oracle.csc_from_coo(src, dst, w, nv)
    return cpu_oracle.katz(offsets, indices, weights, **kw)
