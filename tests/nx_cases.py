"""The reference's Python tests compare cuGraph with NetworkX
(python/cugraph/cugraph/tests/test_pagerank.py:163-267, test_katz_centrality.py:52-113,
test_eigenvector_centrality.py:42-98, test_hits.py:54-124).  The same comparisons, written once
and run by tests/test_python_callers_nx.py (CPU: library replaced by the oracle-backed stand-in,
which makes them an independent check of the ORACLE against NetworkX) and by
tests/test_zz_gpu_python_callers.py (GPU: the CUDA library against NetworkX)."""
import networkx as nx
import numpy as np
import pandas as pd

from tests import util

DATASETS = ["karate.mtx", "dolphins.mtx", "netscience.mtx"]


def edges(name):
    src, dst, w, nv, _ = util.load_mtx(name)
    return pd.DataFrame({"0": src.astype(np.int32), "1": dst.astype(np.int32),
                         "weight": w.astype(np.float32)}), nv


def _host(col):
    return col.cpu().numpy() if hasattr(col, "cpu") else np.asarray(col)


def _sorted_scores(df, column):
    df = df.sort_values("vertex")
    return _host(df["vertex"]), _host(df[column])


def networkx_pagerank(Gnx, max_iter, tol, alpha, personalization_perc, nnz_vtx, seed):
    """test_pagerank.py:99-134 networkx_call."""
    personalization = None
    if personalization_perc != 0:
        rng = np.random.default_rng(seed)
        count = int((nnz_vtx.size * personalization_perc) / 100.0)
        chosen = rng.choice(nnz_vtx, min(nnz_vtx.size, count), replace=False)
        val = rng.random(chosen.size)
        val = val / val.sum()
        personalization = {int(v): float(x) for v, x in zip(chosen, val)}
    z = {k: 1.0 / Gnx.number_of_nodes() for k in Gnx.nodes()}
    pr = nx.pagerank(Gnx, alpha=alpha, nstart=z, max_iter=max_iter * 2, tol=tol * 0.01,
                     personalization=personalization)
    return pr, personalization


def cudify(d, dtype=np.float32):
    """test_pagerank.py:41-48."""
    if d is None:
        return None
    k = np.fromiter(d.keys(), dtype="int32")
    v = np.fromiter(d.values(), dtype=dtype)
    return pd.DataFrame({"vertex": k, "values": v})


def pagerank_case(cugraph, name, personalization_perc, has_guess, has_precomputed, via_nx=False):
    """test_pagerank.py:163-219 (via_nx: :222-267, a NetworkX graph handed straight to
    cugraph.pagerank, which then returns a dict)."""
    max_iter, tol, alpha = 500, 1.0e-06, 0.85
    M, nv = edges(name)
    nnz_vtx = np.unique(M[["0", "1"]])
    Gnx = nx.from_pandas_edgelist(M, source="0", target="1", edge_attr="weight",
                                  create_using=nx.DiGraph())
    networkx_pr, networkx_prsn = networkx_pagerank(Gnx, max_iter, tol, alpha,
                                                   personalization_perc, nnz_vtx, seed=42)
    cu_nstart = pre = None
    if has_guess:
        cu_nstart = cudify(networkx_pr)
        max_iter = 100
    cu_prsn = cudify(networkx_prsn)
    if via_nx:
        pr = cugraph.pagerank(Gnx, alpha=alpha, max_iter=max_iter, tol=tol,
                              personalization=cu_prsn, nstart=cu_nstart)
        assert isinstance(pr, dict)
        vertices = np.array(sorted(pr))
        scores = np.array([pr[v] for v in vertices])
    else:
        G = cugraph.Graph(directed=True)
        G.from_cudf_edgelist(M, source="0", destination="1", edge_attr="weight",
                             store_transposed=True)
        if has_precomputed:
            pre = M[["0", "weight"]].groupby(["0"], as_index=False).sum().rename(
                columns={"0": "vertex", "weight": "sums"})
        df = cugraph.pagerank(G, alpha=alpha, max_iter=max_iter, tol=tol, personalization=cu_prsn,
                              precomputed_vertex_out_weight=pre, nstart=cu_nstart)
        vertices, scores = _sorted_scores(df, "pagerank")
    want = sorted(networkx_pr.items())
    assert len(scores) == len(want)
    err = sum(1 for (v, s), (k, w) in zip(zip(vertices, scores), want)
              if abs(s - w) > tol * 1.1 and v == k)
    assert [k for k, _ in want] == vertices.tolist()
    assert err < 0.01 * len(want), f"{err} mismatches"


def _top_k(vertices, scores, k=10):
    order = np.lexsort((vertices, -np.asarray(scores, np.float64)))
    return vertices[order[:k]].tolist()


def katz_case(cugraph, name):
    """test_katz_centrality.py:52-82: unweighted directed graph, alpha = 1 / max degree, the ten
    best vertices agree with NetworkX."""
    M, nv = edges(name)
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist(M, source="0", destination="1", store_transposed=True)
    katz_alpha = 1 / int(G.degree()["degree"].max())
    vertices, cu = _sorted_scores(cugraph.katz_centrality(G, alpha=None, max_iter=1000),
                                  "katz_centrality")
    Gnx = nx.from_pandas_edgelist(M, create_using=nx.DiGraph(), source="0", target="1")
    nk = nx.katz_centrality(Gnx, alpha=katz_alpha)
    ref = np.array([nk[k] for k in sorted(nk.keys())])
    assert sorted(nk.keys()) == vertices.tolist()
    assert _top_k(vertices, cu) == _top_k(vertices, ref)
    # both are L2-normalised: the scores themselves agree too
    assert np.abs(cu - ref).max() <= 1e-4


def eigenvector_case(cugraph, name):
    """test_eigenvector_centrality.py:42-69."""
    M, nv = edges(name)
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist(M, source="0", destination="1", store_transposed=True)
    vertices, cu = _sorted_scores(cugraph.eigenvector_centrality(G, max_iter=1000),
                                  "eigenvector_centrality")
    Gnx = nx.from_pandas_edgelist(M, create_using=nx.DiGraph(), source="0", target="1")
    nk = nx.eigenvector_centrality(Gnx, max_iter=1000)
    ref = np.array([nk[k] for k in sorted(nk.keys())])
    assert _top_k(vertices, cu) == _top_k(vertices, ref)
    assert np.abs(cu - ref).max() <= 1e-3


def hits_case(cugraph, name, max_iter=50, tol=1.0e-6):
    """test_hits.py:54-124: hubs within 1e-5 and authorities within 1e-4 of nx.hits."""
    M, nv = edges(name)
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist(M, source="0", destination="1", edge_attr="weight",
                         store_transposed=True)
    df = cugraph.hits(G, max_iter, tol).sort_values("vertex")
    # HITS in the library ignores edge weights (hits_impl.cuh:114-140); the reference test's
    # datasets have unit weights, netscience does not: compare with the unweighted NetworkX graph
    Gnx = nx.from_pandas_edgelist(M, create_using=nx.DiGraph(), source="0", target="1")
    nx_hubs, nx_authorities = nx.hits(Gnx, max_iter, tol, normalized=True)
    keys = sorted(nx_hubs)
    assert keys == _host(df["vertex"]).tolist()
    assert np.abs(_host(df["hubs"]) - np.array([nx_hubs[k] for k in keys])).max() <= 0.00001
    assert np.abs(_host(df["authorities"]) -
                  np.array([nx_authorities[k] for k in keys])).max() <= 0.0001


def nx_input_case(cugraph):
    """NetworkX graphs as inputs (utilities/nx_factory.py convert_from_nx): undirected karate
    club with string labels; results come back as dicts keyed by the labels."""
    Gnx = nx.relabel_nodes(nx.karate_club_graph(), lambda n: f"m{n:02d}")
    for _, _, data in Gnx.edges(data=True):
        data.clear()  # unweighted
    pr = cugraph.pagerank(Gnx, tol=1e-6, max_iter=500)
    want = nx.pagerank(Gnx, alpha=0.85, tol=1e-8, max_iter=1000)
    assert sorted(pr) == sorted(want)
    assert max(abs(pr[k] - want[k]) for k in want) <= 2e-6
    kz = cugraph.katz_centrality(Gnx, alpha=0.02, tol=1e-6, max_iter=1000)
    kw = nx.katz_centrality(Gnx, alpha=0.02, tol=1e-9, max_iter=5000)
    assert max(abs(kz[k] - kw[k]) for k in kw) <= 1e-4
    ev = cugraph.eigenvector_centrality(Gnx, tol=1e-6, max_iter=1000)
    ew = nx.eigenvector_centrality(Gnx, tol=1e-9, max_iter=5000)
    assert max(abs(ev[k] - ew[k]) for k in ew) <= 1e-3
    hubs, auth = cugraph.hits(Gnx, max_iter=100, tol=1e-6)
    nh, na = nx.hits(Gnx, max_iter=500, tol=1e-10, normalized=True)
    assert max(abs(hubs[k] - nh[k]) for k in nh) <= 1e-4
    assert max(abs(auth[k] - na[k]) for k in na) <= 1e-4
    # a weighted DiGraph (nx.is_weighted looks for the attribute called "weight", as in the
    # reference): by name, and as the graph's only edge attribute when weight=None
    D = nx.DiGraph()
    D.add_weighted_edges_from([(0, 1, 0.1), (1, 3, 2.1), (1, 4, 1.1), (2, 0, 5.1), (2, 1, 3.1),
                               (2, 3, 4.1), (3, 5, 7.2), (4, 5, 3.2)])
    want = [0.0915528, 0.168382, 0.0656831, 0.191468, 0.120677, 0.362237]  # pagerank_test.c
    for weight in ("weight", None):
        pr = cugraph.pagerank(D, alpha=0.95, tol=1e-4, max_iter=20, weight=weight)
        assert all(abs(pr[i] - want[i]) <= 1e-3 * max(pr[i], want[i]) for i in range(6))
