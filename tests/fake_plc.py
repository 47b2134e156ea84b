"""TEST INFRASTRUCTURE ONLY: an oracle-backed stand-in for cugraph_forked_b200.pylibcugraph.

Lets the CPU suite drive the cugraph-shaped Python callers (cugraph_forked_b200/cugraph.py:
table handling, symmetrize, renumbering, argument checks) end to end without a GPU.  Same
function names and keyword arguments as the real module; the arithmetic is the CPU oracle
(oracle/cpu_oracle.py).  Like the library, a graph made with renumber=True hands results back
under the caller's vertex labels, in no particular order (a fixed shuffle here, so that callers
cannot rely on one)."""
import numpy as np
import torch

from oracle import cpu_oracle


class ResourceHandle:
    def __init__(self, handle=None):
        pass


class GraphProperties:
    def __init__(self, is_symmetric=False, is_multigraph=False):
        self.is_symmetric = is_symmetric
        self.is_multigraph = is_multigraph


class SGGraph:
    def __init__(self, resource_handle, graph_properties, src_array, dst_array, weight_array,
                 store_transposed=False, renumber=False, do_expensive_check=False,
                 edge_id_array=None, edge_type_array=None):
        for a in (src_array, dst_array):
            assert isinstance(a, torch.Tensor) and a.is_contiguous()
        if src_array.dtype != dst_array.dtype:
            raise ValueError("CUGRAPH_INVALID_INPUT src type != dst type")
        self.vertex_dtype = src_array.dtype
        src, dst = src_array.numpy(), dst_array.numpy()
        self.props = graph_properties
        self.store_transposed = store_transposed
        if renumber:
            self.labels = np.unique(np.concatenate([src, dst]))
            src, dst = np.searchsorted(self.labels, src), np.searchsorted(self.labels, dst)
        else:
            nv = int(max(src.max(initial=-1), dst.max(initial=-1))) + 1
            self.labels = np.arange(nv, dtype=src.dtype)
        self.nv = len(self.labels)
        self.src, self.dst = src.astype(np.int32), dst.astype(np.int32)
        self.w = None if weight_array is None else weight_array.numpy()
        if self.w is not None and self.w.dtype not in (np.float32, np.float64):
            raise ValueError("CUGRAPH_UNSUPPORTED_TYPE_COMBINATION weight type")
        self.dtype = np.float32 if self.w is None else self.w.dtype.type

    def internal(self, vertices):
        assert vertices.dtype == self.vertex_dtype, "vertex id type differs from the graph's"
        v = vertices.numpy()
        pos = np.searchsorted(self.labels, v)
        ok = (pos < self.nv) & (self.labels[np.minimum(pos, self.nv - 1)] == v)
        if not ok.all():
            raise ValueError("CUGRAPH_INVALID_INPUT vertex is not in the graph")
        return pos

    def dense(self, vertices, values, fill=0.0):
        out = np.full(self.nv, fill, self.dtype)
        assert values.numpy().dtype == self.dtype, "value type must match the graph's"
        out[self.internal(vertices)] = values.numpy()
        return out

    def result(self, *columns):
        perm = np.random.default_rng(7).permutation(self.nv)
        return (torch.from_numpy(self.labels[perm].copy()),) + tuple(
            torch.from_numpy(np.ascontiguousarray(c[perm])) for c in columns)


def _pagerank(graph, pre_v, pre_s, guess_v, guess_vals, pers, alpha, epsilon, max_iterations):
    offsets, indices, weights = cpu_oracle.csc_from_coo(graph.src, graph.dst, graph.w, graph.nv)
    if pre_v is not None:
        # the oracle derives the out-weight sums itself; check the caller's agree
        ow = np.zeros(graph.nv)
        np.add.at(ow, graph.src, 1.0 if graph.w is None else graph.w)
        given = graph.dense(pre_v, pre_s)
        idx = graph.internal(pre_v)
        np.testing.assert_allclose(given[idx], ow[idx], rtol=1e-5)
    guess = None if guess_v is None else graph.dense(guess_v, guess_vals)
    personalization = None
    if pers is not None:
        personalization = (graph.internal(pers[0]).astype(np.int32), pers[1].numpy())
        assert pers[1].numpy().dtype == graph.dtype
    scores, _ = cpu_oracle.pagerank(offsets, indices, weights, dtype=graph.dtype, alpha=alpha,
                                    epsilon=epsilon, max_iterations=max_iterations,
                                    personalization=personalization, initial_guess=guess)
    return graph.result(scores)


def pagerank(resource_handle, graph, precomputed_vertex_out_weight_vertices,
             precomputed_vertex_out_weight_sums, initial_guess_vertices, initial_guess_values,
             alpha, epsilon, max_iterations, do_expensive_check):
    return _pagerank(graph, precomputed_vertex_out_weight_vertices,
                     precomputed_vertex_out_weight_sums, initial_guess_vertices,
                     initial_guess_values, None, alpha, epsilon, max_iterations)


def personalized_pagerank(resource_handle, graph, precomputed_vertex_out_weight_vertices,
                          precomputed_vertex_out_weight_sums, initial_guess_vertices,
                          initial_guess_values, personalization_vertices, personalization_values,
                          alpha, epsilon, max_iterations, do_expensive_check):
    return _pagerank(graph, precomputed_vertex_out_weight_vertices,
                     precomputed_vertex_out_weight_sums, initial_guess_vertices,
                     initial_guess_values, (personalization_vertices, personalization_values),
                     alpha, epsilon, max_iterations)


def katz_centrality(resource_handle, graph, betas, alpha, beta, epsilon, max_iterations,
                    do_expensive_check):
    offsets, indices, weights = cpu_oracle.csc_from_coo(graph.src, graph.dst, graph.w, graph.nv)
    scores, _ = cpu_oracle.katz(offsets, indices, weights, dtype=graph.dtype, alpha=alpha,
                                beta=beta, epsilon=epsilon, max_iterations=max_iterations,
                                normalize=True)
    return graph.result(scores)


def eigenvector_centrality(resource_handle, graph, epsilon, max_iterations, do_expensive_check):
    # the reference oracle pulls into tmp[src[e]] from old[dst[e]]: feed it the transposed list
    scores, _ = cpu_oracle.eigenvector(graph.dst, graph.src, graph.w, num_vertices=graph.nv,
                                       dtype=graph.dtype, epsilon=epsilon,
                                       max_iterations=max_iterations)
    return graph.result(scores)


def hits(resource_handle, graph, tol, max_iter, initial_hubs_guess_vertices,
         initial_hubs_guess_values, normalized, do_expensive_check):
    offsets, indices = cpu_oracle.csr_from_coo(graph.src, graph.dst, graph.nv)
    start = None
    if initial_hubs_guess_vertices is not None:
        start = graph.dense(initial_hubs_guess_vertices, initial_hubs_guess_values)
    hubs, auth, _, _ = cpu_oracle.hits(offsets, indices, dtype=graph.dtype, max_iterations=max_iter,
                                       tolerance=tol, starting_hubs=start, normalize=normalized)
    return graph.result(hubs, auth)


class MGGraph(SGGraph):
    """World size 1 only: the whole edge list is this rank's slice; renumbered by the library."""

    def __init__(self, resource_handle, graph_properties, src_array, dst_array, weight_array,
                 store_transposed=False, num_edges=-1, do_expensive_check=False,
                 edge_id_array=None, edge_type_array=None):
        super().__init__(resource_handle, graph_properties, src_array, dst_array, weight_array,
                         store_transposed=store_transposed, renumber=True)
