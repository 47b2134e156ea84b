"""CPU: the cugraph-shaped Python callers (cugraph_forked_b200/cugraph.py) - tables, symmetrize,
renumbering, argument checks and warnings - with the CUDA library replaced by the oracle-backed
stand-in tests/fake_plc.py.  Modelled on the reference's python/cugraph/cugraph/tests/
test_pagerank.py, test_katz_centrality.py, test_eigenvector_centrality.py, test_hits.py,
test_graph.py, test_symmetrize.py and test_renumber.py."""
import numpy as np
import pandas as pd
import pytest
import torch

from tests import fake_plc
from tests import golden_vectors as gv
from tests import util


@pytest.fixture()
def cugraph(monkeypatch, oracle_built):
    from cugraph_forked_b200 import cugraph as cg
    monkeypatch.setattr(cg, "_plc", fake_plc)
    monkeypatch.setattr(cg, "_DEVICE", "cpu")
    return cg


def _karate():
    src, dst, w, nv, _ = util.load_mtx("karate.mtx")
    assert nv == 34
    return src.astype(np.int32), dst.astype(np.int32), w.astype(np.float32)


def _scores(df, column, nv, key="vertex"):
    v = df[key].numpy()
    assert sorted(v.tolist()) == list(range(nv))
    out = np.empty(nv, df[column].numpy().dtype)
    out[v] = df[column].numpy()
    return out


# ---- Frame ---------------------------------------------------------------------------------

def test_frame_columns_and_rows(cugraph):
    f = cugraph.Frame({"a": [3, 1, 2], "b": np.array([0.5, 1.5, 2.5], np.float32)})
    assert f.columns == ["a", "b"] and len(f) == 3
    assert f["a"].dtype == torch.int64 and f["b"].dtype == torch.float32
    assert f[["b"]].columns == ["b"]
    assert f[f["a"] > 1]["b"].tolist() == [0.5, 2.5]
    assert f.sort_values("a")["b"].tolist() == [1.5, 2.5, 0.5]
    assert f.rename({"a": "x"}).columns == ["x", "b"]
    assert f.drop("a").columns == ["b"]
    assert f.astype({"b": np.float64})["b"].dtype == torch.float64
    pdf = f.to_pandas()
    assert list(pdf.columns) == ["a", "b"] and pdf["a"].tolist() == [3, 1, 2]
    with pytest.raises(ValueError, match="rows"):
        f["c"] = [1, 2]
    g = cugraph.Frame(pd.DataFrame({"s": ["x", "y"], "v": [1, 2]}))
    assert isinstance(g["s"], np.ndarray) and g.sort_values("s")["v"].tolist() == [1, 2]
    with pytest.raises(ValueError, match="NULL"):
        cugraph.Frame(pd.DataFrame({"v": [1.0, None]}))


# ---- symmetrize ----------------------------------------------------------------------------

@pytest.mark.parametrize("do_symmetrize", [True, False])
@pytest.mark.parametrize("weighted", [True, False])
def test_symmetrize_matches_groupby_min(cugraph, do_symmetrize, weighted):
    """structure/symmetrize.py symmetrize_df: concat with the mirrored table, groupby min."""
    rng = np.random.default_rng(3)
    src = rng.integers(0, 40, 3000)
    dst = rng.integers(0, 40, 3000)
    w = rng.random(3000).astype(np.float32)
    df = pd.DataFrame({"s": src, "d": dst, "w": w})
    if do_symmetrize:
        df = pd.concat([df, df.rename(columns={"s": "d", "d": "s"})[["s", "d", "w"]]])
    want = df.groupby(["s", "d"], as_index=False).min()
    s, d, vals = cugraph.symmetrize(torch.from_numpy(src), torch.from_numpy(dst),
                                    [torch.from_numpy(w)] if weighted else None,
                                    symmetrize=do_symmetrize)
    assert s.tolist() == want["s"].tolist() and d.tolist() == want["d"].tolist()
    if weighted:
        assert np.array_equal(vals[0].numpy(), want["w"].to_numpy())
    else:
        assert vals is None
    # a MultiGraph keeps the parallel edges
    s2, _, _ = cugraph.symmetrize(torch.from_numpy(src), torch.from_numpy(dst), None, multi=True,
                                  symmetrize=do_symmetrize)
    assert s2.numel() == (6000 if do_symmetrize else 3000)


def test_symmetrize_empty(cugraph):
    e = torch.zeros(0, dtype=torch.int32)
    s, d, v = cugraph.symmetrize(e, e, [torch.zeros(0)])
    assert s.numel() == 0 and d.numel() == 0 and v[0].numel() == 0


# ---- NumberMap -----------------------------------------------------------------------------

def test_number_map_integers(cugraph):
    src = torch.tensor([100, 7, 7, 42], dtype=torch.int64)
    dst = torch.tensor([42, 100, 9, 9], dtype=torch.int64)
    m = cugraph.NumberMap([src], [dst], ["s"])
    assert m.num_vertices == 4 and m.id_type == torch.int32
    ids = m.to_internal_vertex_id([torch.tensor([7, 9, 42, 100, 8, -5, 1000])])
    assert ids.tolist() == [0, 1, 2, 3, -1, -1, -1]
    assert m.from_internal_vertex_id(torch.tensor([3, 0]))[0].tolist() == [100, 7]


def test_number_map_strings_and_multi_column(cugraph):
    m = cugraph.NumberMap([np.array(["b", "a", "c"])], [np.array(["a", "d", "b"])], ["s"])
    assert m.to_internal_vertex_id([np.array(["a", "b", "c", "d", "zz"])]).tolist() == \
        [0, 1, 2, 3, -1]
    assert m.from_internal_vertex_id(torch.tensor([3, 1]))[0].tolist() == ["d", "b"]
    s0, s1 = torch.tensor([5, 1, 5]), torch.tensor([1005, 1001, 1005])
    d0, d1 = torch.tensor([1, 9, 9]), torch.tensor([1001, 1009, 1009])
    m2 = cugraph.NumberMap([s0, s1], [d0, d1], ["s0", "s1"])
    assert m2.num_vertices == 3 and m2.vertex_column_size() == 2
    ids = m2.to_internal_vertex_id([torch.tensor([9, 1, 5, 5]), torch.tensor([1009, 1001, 1005, 0])])
    assert ids.tolist() == [2, 0, 1, -1]
    df = m2.unrenumber({"vertex": torch.tensor([2, 0]), "x": torch.tensor([0.5, 0.25])}, "vertex")
    assert df.columns == ["0_vertex", "1_vertex", "x"]
    assert df["0_vertex"].tolist() == [9, 1] and df["1_vertex"].tolist() == [1009, 1001]
    with pytest.raises(ValueError, match="vertex column"):
        m2.to_internal_vertex_id([torch.tensor([1])])


# ---- Graph ---------------------------------------------------------------------------------

def test_graph_directed_properties(cugraph):
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist({"s": gv.G6_SRC, "d": gv.G6_DST, "w": gv.G6_WGT}, "s", "d", "w",
                         store_transposed=True)
    assert G.is_directed() and G.is_weighted() and not G.is_renumbered()
    assert not G.is_multigraph() and G.store_transposed is True
    assert G.number_of_vertices() == 6 == G.number_of_nodes()
    assert G.number_of_edges() == 8
    assert G.in_degree()["degree"].tolist() == np.bincount(gv.G6_DST, minlength=6).tolist()
    assert G.out_degree()["degree"].tolist() == np.bincount(gv.G6_SRC, minlength=6).tolist()
    both = np.bincount(gv.G6_DST, minlength=6) + np.bincount(gv.G6_SRC, minlength=6)
    assert G.degree()["degree"].tolist() == both.tolist()
    sub = G.degrees([1, 5])
    assert sub.columns == ["vertex", "in_degree", "out_degree"]
    assert sub["vertex"].tolist() == [1, 5] and sub["in_degree"].tolist() == [2, 2]
    el = G.view_edge_list()
    assert el.columns == ["src", "dst", "weights"] and len(el) == 8
    assert G.has_node(3) and not G.has_node(17)
    assert G.has_edge(2, 3) and not G.has_edge(3, 2)
    assert sorted(G.nodes().tolist()) == list(range(6))
    with pytest.raises(RuntimeError, match="Graph already has values"):
        G.from_cudf_edgelist({"s": gv.G6_SRC, "d": gv.G6_DST}, "s", "d")
    G.clear()
    with pytest.raises(RuntimeError, match="Graph is Empty"):
        G.number_of_vertices()


def test_graph_undirected_symmetrizes_and_drops_duplicates(cugraph):
    G = cugraph.Graph()
    df = pd.DataFrame({"source": [0, 1, 1, 2, 0], "destination": [1, 0, 2, 1, 1],
                       "w": [5.0, 3.0, 2.0, 7.0, 4.0]})
    G.from_pandas_edgelist(df, edge_attr="w")
    el = G.edgelist.edgelist_df
    assert list(zip(el["src"].tolist(), el["dst"].tolist())) == [(0, 1), (1, 0), (1, 2), (2, 1)]
    assert el["weights"].tolist() == [3.0, 3.0, 2.0, 2.0]
    assert G.number_of_edges() == 2 and G.number_of_edges(directed_edges=True) == 4
    assert G.degree()["degree"].tolist() == [2, 4, 2]
    view = G.view_edge_list()
    assert list(zip(view["src"].tolist(), view["dst"].tolist())) == [(0, 1), (1, 2)]
    D = G.to_directed()
    assert D.is_directed() and D.number_of_edges() == 4
    U = D.to_undirected()
    assert not U.is_directed() and U.number_of_edges() == 2
    M = cugraph.MultiGraph(directed=True)
    M.from_pandas_edgelist(df, edge_attr="w")
    assert M.is_multigraph() and M.number_of_edges() == 5
    G2 = cugraph.Graph(M, directed=True)
    assert G2.number_of_edges() == 4


def test_graph_input_errors(cugraph):
    df = {"s": gv.G6_SRC, "d": gv.G6_DST, "w": gv.G6_WGT}
    G = cugraph.Graph(directed=True)
    with pytest.raises(ValueError, match="source column names and/or destination column"):
        G.from_cudf_edgelist(df, "s", "nope")
    with pytest.raises(ValueError, match="edge_attr column name not found"):
        G.from_cudf_edgelist(df, "s", "d", "nope")
    with pytest.raises(ValueError, match="Invalid number of edge attributes"):
        G.from_cudf_edgelist(df, "s", "d", ["w", "w"])
    with pytest.raises(ValueError, match="not permitted for an undirected graph"):
        cugraph.Graph().from_cudf_edgelist(df, "s", "d", ["w", "s", "d"])
    with pytest.raises(TypeError, match="input should be"):
        G.from_cudf_edgelist([1, 2, 3], "s", "d")
    with pytest.raises(TypeError, match="not a Pandas DataFrame"):
        G.from_pandas_edgelist(df, "s", "d")
    strings = {"s": np.array(["a", "b"]), "d": np.array(["b", "c"])}
    with pytest.raises(ValueError, match="set renumber to True for non integer"):
        G.from_cudf_edgelist(strings, "s", "d", renumber=False)
    multi = {"s0": [0, 1], "s1": [5, 6], "d0": [1, 0], "d1": [6, 5]}
    with pytest.raises(ValueError, match="set renumber to True for multi column"):
        G.from_cudf_edgelist(multi, ["s0", "s1"], ["d0", "d1"], renumber=False)
    with pytest.raises(ValueError, match="NULL"):
        G.from_cudf_edgelist(pd.DataFrame({"s": [0.0, None], "d": [1.0, 2.0]}), "s", "d")
    assert G.edgelist is None  # nothing was loaded by the failed calls


def test_integer_weights_are_widened(cugraph):
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist({"s": gv.G6_SRC, "d": gv.G6_DST, "w": np.arange(8, dtype=np.int32) + 1},
                         "s", "d", "w")
    assert G._plc_graph.dtype == np.float32
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist({"s": gv.G6_SRC, "d": gv.G6_DST, "w": np.arange(8, dtype=np.int64) + 1},
                         "s", "d", "w")
    assert G._plc_graph.dtype == np.float64


def test_from_numpy_array(cugraph):
    a = np.zeros((4, 4))
    a[0, 1] = a[1, 0] = 2.0
    a[2, 3] = a[3, 2] = 1.5
    G = cugraph.Graph()
    G.from_numpy_array(a)
    assert G.number_of_edges() == 2 and G.is_weighted()
    with pytest.raises(TypeError, match="not a Numpy array"):
        cugraph.Graph().from_numpy_array([[0, 1], [1, 0]])


def test_adjacency_list_round_trip(cugraph):
    """graph_classes.py:173 from_cudf_adjlist and simpleGraph.py:372,416 view_(transposed_)adj_list
    against scipy's CSR / CSC of the same matrix."""
    import scipy.sparse as sp
    src, dst, w = _karate()
    M = sp.coo_matrix((w, (src, dst)), shape=(34, 34)).tocsr()
    M.sort_indices()
    G = cugraph.Graph(directed=True)
    G.from_cudf_adjlist(M.indptr, M.indices, M.data, store_transposed=True)
    assert G.number_of_edges() == len(src) and G.is_weighted()
    off, ind, val = G.view_adj_list()
    assert np.array_equal(off.numpy(), M.indptr) and np.array_equal(ind.numpy(), M.indices)
    assert np.array_equal(val.numpy(), M.data)
    T = M.tocsc()
    T.sort_indices()
    off, ind, val = G.view_transposed_adj_list()
    assert np.array_equal(off.numpy(), T.indptr) and np.array_equal(ind.numpy(), T.indices)
    assert np.array_equal(val.numpy(), T.data)
    want, _ = util.oracle_pagerank(src, dst, w, 34, dtype=np.float32, alpha=0.85, epsilon=1e-6,
                                   max_iterations=500)
    got = _scores(cugraph.pagerank(G, tol=1e-6, max_iter=500), "pagerank", 34)
    assert np.abs(got - want).max() <= 2e-6
    U = cugraph.Graph(directed=True)
    U.from_cudf_adjlist(M.indptr, M.indices)
    assert not U.is_weighted() and U.view_adj_list()[2] is None
    with pytest.raises(ValueError, match="run from 0 to len"):
        cugraph.Graph().from_cudf_adjlist(M.indptr[:-1], M.indices)
    with pytest.raises(ValueError, match="non-decreasing"):
        cugraph.Graph().from_cudf_adjlist(np.array([0, 3, 2, 4]), np.array([1, 2, 0, 1]))
    with pytest.raises(RuntimeError, match="already has values"):
        G.from_cudf_adjlist(M.indptr, M.indices)


def test_distributed_loader_world_size_1(cugraph):
    """from_distributed_edgelist (the from_dask_cudf_edgelist stand-in) with one rank holding
    the whole list: same scores as the single-GPU loader; undirected slices are mirrored."""
    src, dst, w = _karate()
    rh = fake_plc.ResourceHandle()
    G = cugraph.Graph(directed=True)
    G.from_distributed_edgelist({"s": src, "d": dst, "w": w}, "s", "d", "w", resource_handle=rh,
                                store_transposed=True, num_global_edges=len(src))
    assert G.distributed and G._handle is rh
    want, _ = util.oracle_pagerank(src, dst, w, 34, dtype=np.float32, alpha=0.85, epsilon=1e-6,
                                   max_iterations=500)
    got = _scores(cugraph.pagerank(G, tol=1e-6, max_iter=500), "pagerank", 34)
    assert np.abs(got - want).max() <= 2e-6
    with pytest.raises(NotImplementedError, match="only its slice"):
        G.number_of_vertices()
    with pytest.raises(NotImplementedError, match="only its slice"):
        cugraph.katz_centrality(G)  # alpha=None needs the global degrees
    kz = _scores(cugraph.katz_centrality(G, alpha=0.02, tol=1e-6, max_iter=500),
                 "katz_centrality", 34)
    kw, _ = util.oracle_katz(src, dst, w, 34, dtype=np.float32, alpha=0.02, beta=1.0, epsilon=1e-6,
                             max_iterations=500, normalize=True)
    assert np.abs(kz - kw).max() <= 1e-6
    keep = src < dst
    U = cugraph.Graph()
    U.from_distributed_edgelist({"s": src[keep], "d": dst[keep]}, "s", "d", resource_handle=rh)
    assert len(U.edgelist.edgelist_df) == 2 * int(keep.sum())
    with pytest.raises(ValueError, match="needs the multi-GPU resource_handle"):
        cugraph.Graph().from_distributed_edgelist({"s": src, "d": dst}, "s", "d")
    with pytest.raises(ValueError, match="single integer id columns"):
        cugraph.Graph().from_distributed_edgelist({"s": np.array(["a"]), "d": np.array(["b"])},
                                                  "s", "d", resource_handle=rh)


# ---- pagerank ------------------------------------------------------------------------------

def _toy(cugraph, store_transposed=True, dtype=np.float32):
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist({"s": gv.G6_SRC, "d": gv.G6_DST, "w": gv.G6_WGT.astype(dtype)}, "s", "d",
                         "w", store_transposed=store_transposed)
    return G


def test_pagerank_toy_golden(cugraph):
    """The C-API golden vector (c_api/pagerank_test.c) through the Python caller."""
    p = gv.G6_PAGERANK
    df = cugraph.pagerank(_toy(cugraph), alpha=p["alpha"], tol=p["epsilon"],
                          max_iter=p["max_iterations"])
    assert df.columns == ["vertex", "pagerank"]
    assert gv.nearly_equal(_scores(df, "pagerank", 6), p["result"])


@pytest.mark.parametrize("has_guess", [False, True])
@pytest.mark.parametrize("has_precomputed", [False, True])
@pytest.mark.parametrize("personalized", [False, True])
def test_pagerank_karate(cugraph, has_guess, has_precomputed, personalized):
    """test_pagerank.py:163-219 with the oracle where the reference uses NetworkX."""
    src, dst, w = _karate()
    nv = 34
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist({"0": src, "1": dst, "2": w}, "0", "1", "2", store_transposed=True)
    pers = None
    kw = {}
    if personalized:
        pv = np.array([3, 17, 26, 30], np.int32)
        pw = np.array([0.4, 0.3, 0.2, 0.1], np.float32)
        pers = {"vertex": pv, "values": pw}
        kw["personalization"] = (pv, pw)
    want, _ = util.oracle_pagerank(src, dst, w, nv, dtype=np.float32, alpha=0.85, epsilon=1e-6,
                                   max_iterations=500, **kw)
    nstart = pre = None
    if has_guess:
        perm = np.random.default_rng(0).permutation(nv).astype(np.int32)
        nstart = {"vertex": perm, "values": want[perm]}
    if has_precomputed:
        ow = np.zeros(nv, np.float32)
        np.add.at(ow, src, w)
        pre = pd.DataFrame({"vertex": np.arange(nv, dtype=np.int32), "sums": ow})
    df = cugraph.pagerank(G, alpha=0.85, personalization=pers, precomputed_vertex_out_weight=pre,
                          max_iter=500, tol=1e-6, nstart=nstart)
    got = _scores(df, "pagerank", nv)
    assert np.abs(got - want).max() <= 2e-6


def test_pagerank_string_labels(cugraph):
    """Non-integer vertex labels go through the NumberMap and come back as labels."""
    src, dst, w = _karate()
    names = np.array([f"v{i:02d}" for i in range(34)])
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist(pd.DataFrame({"a": names[src], "b": names[dst], "w": w}), "a", "b", "w",
                         store_transposed=True)
    assert G.is_renumbered() and G.number_of_vertices() == 34
    pers = pd.DataFrame({"vertex": ["v17", "v03"], "values": np.array([0.5, 0.5], np.float32)})
    df = cugraph.pagerank(G, personalization=pers, tol=1e-6, max_iter=500)
    want, _ = util.oracle_pagerank(src, dst, w, 34, dtype=np.float32, alpha=0.85, epsilon=1e-6,
                                   max_iterations=500,
                                   personalization=(np.array([17, 3]), np.array([0.5, 0.5])))
    got = df.sort_values("vertex")
    assert got["vertex"].tolist() == names.tolist()
    assert np.abs(got["pagerank"].numpy() - want).max() <= 2e-6
    with pytest.raises(ValueError, match="not in the graph"):
        cugraph.pagerank(G, personalization=pd.DataFrame(
            {"vertex": ["nobody"], "values": np.array([1.0], np.float32)}))
    assert sorted(G.nodes().tolist()) == names.tolist()
    assert G.has_node("v05") and not G.has_node("v99")
    assert G.lookup_internal_vertex_id(np.array(["v02", "zz"])).tolist() == [2, -1]


@pytest.mark.parametrize("has_guess", [False, True])
@pytest.mark.parametrize("has_precomputed", [False, True])
def test_pagerank_multi_column(cugraph, has_guess, has_precomputed):
    """test_pagerank.py:270-362: two-column vertex labels; input tables carry the label columns
    first and the value column last; the result has '0_vertex' and '1_vertex'."""
    src, dst, w = _karate()
    nv = 34
    want, _ = util.oracle_pagerank(src, dst, w, nv, dtype=np.float32, alpha=0.85, epsilon=1e-6,
                                   max_iterations=500,
                                   personalization=(np.array([4, 9]), np.array([0.5, 0.5])))
    cu_M = pd.DataFrame({"src_0": src, "dst_0": dst, "src_1": src + 1000, "dst_1": dst + 1000,
                         "weights": w})
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist(cu_M, source=["src_0", "src_1"], destination=["dst_0", "dst_1"],
                         edge_attr="weights", store_transposed=True)
    prsn = pd.DataFrame({"vertex_0": np.array([4, 9], np.int32),
                         "vertex_1": np.array([1004, 1009], np.int32),
                         "values": np.array([0.5, 0.5], np.float32)})
    nstart = pre = None
    if has_guess:
        v = np.arange(nv, dtype=np.int32)
        nstart = pd.DataFrame({"vertex_0": v, "vertex_1": v + 1000, "values": want})
    if has_precomputed:
        pre = cu_M[["src_0", "src_1", "weights"]].groupby(
            ["src_0", "src_1"], as_index=False).sum().rename(columns={"weights": "sums"})
    df = cugraph.pagerank(G, alpha=0.85, max_iter=500, tol=1e-6, personalization=prsn,
                          nstart=nstart, precomputed_vertex_out_weight=pre)
    assert df.columns == ["0_vertex", "1_vertex", "pagerank"]
    df = df.sort_values("0_vertex")
    assert df["0_vertex"].tolist() == list(range(nv))
    assert df["1_vertex"].tolist() == [i + 1000 for i in range(nv)]
    assert np.abs(df["pagerank"].numpy() - want).max() <= 2e-6


def test_pagerank_unweighted_and_sparse_ids(cugraph):
    """An unweighted edge list computes in float32 with unit weights; integer labels with gaps
    are renumbered by the library and only the labels that occur come back."""
    rng = np.random.default_rng(11)
    lab = np.sort(rng.choice(100000, 200, replace=False)).astype(np.int64)
    s, d = rng.integers(0, 200, 1500), rng.integers(0, 200, 1500)
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist({"source": lab[s], "destination": lab[d]}, store_transposed=True)
    assert not G.is_weighted() and not G.is_renumbered()
    df = cugraph.pagerank(G, tol=1e-6, max_iter=500)
    assert df["pagerank"].dtype == torch.float32 and df["vertex"].dtype == torch.int64
    present = np.unique(np.concatenate([s, d]))
    assert sorted(df["vertex"].tolist()) == lab[present].tolist()
    pairs = np.unique(np.stack([s, d]), axis=1)  # the Graph drops duplicate edges
    dense = np.searchsorted(present, pairs)
    want, _ = util.oracle_pagerank(dense[0], dense[1], None, len(present), dtype=np.float32,
                                   alpha=0.85, epsilon=1e-6, max_iterations=500)
    got = df.sort_values("vertex")["pagerank"].numpy()
    assert np.abs(got - want).max() <= 2e-6
    assert abs(float(got.sum()) - 1.0) < 1e-4


def test_pagerank_renumber_false_keeps_isolated_ids(cugraph):
    """pagerank.py docstring NOTE: with renumber=False ids are taken as 0..max and the missing
    ones are isolated vertices that still get a score."""
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist({"s": np.array([0, 5], np.int32), "d": np.array([5, 2], np.int32)}, "s",
                         "d", renumber=False, store_transposed=True)
    df = cugraph.pagerank(G)
    assert sorted(df["vertex"].tolist()) == list(range(6))


def test_pagerank_invalid_personalization_dtype(cugraph):
    """test_pagerank.py:365-388, same warning text."""
    src, dst, w = _karate()
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist({"src": src, "dst": dst, "weights": w}, source="src", destination="dst",
                         edge_attr="weights", store_transposed=True)
    personalization_vec = pd.DataFrame({"vertex": [17, 26], "values": [0.5, 0.75]})
    warning_msg = ("PageRank requires 'personalization' values to match the "
                   "graph's 'edge_attr' type. edge_attr type is: "
                   "float32 and got 'personalization' values "
                   "of type: float64.")
    with pytest.warns(UserWarning, match=warning_msg):
        cugraph.pagerank(G, personalization=personalization_vec)
    with pytest.warns(UserWarning, match="'nstart' values"):
        cugraph.pagerank(G, nstart={"vertex": np.arange(34), "values": np.full(34, 1 / 34)})
    with pytest.raises(NotImplementedError, match="personalization other than"):
        cugraph.pagerank(G, personalization=[0.5, 0.5])


def test_transposed_false_warnings(cugraph):
    """test_pagerank.py:391, test_katz_centrality.py:176, test_eigenvector_centrality.py:156,
    test_hits.py:127."""
    src, dst, w = _karate()
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist({"0": src, "1": dst, "2": w}, "0", "1", "2")
    for fn, name in ((cugraph.pagerank, "Pagerank"), (cugraph.katz_centrality, "Katz centrality"),
                     (cugraph.eigenvector_centrality, "Eigenvector centrality"),
                     (cugraph.hits, "HITS")):
        msg = (f"{name} expects the 'store_transposed' flag to be set to 'True' for optimal "
               "performance during the graph creation")
        with pytest.warns(UserWarning, match=msg):
            fn(G)


def test_algorithms_need_a_loaded_graph(cugraph):
    with pytest.raises(RuntimeError, match="Graph is Empty"):
        cugraph.pagerank(cugraph.Graph(directed=True))
    with pytest.raises(TypeError, match="input must be either a cuGraph or NetworkX graph"):
        cugraph.pagerank({"not": "a graph"})


# ---- katz / eigenvector / hits -------------------------------------------------------------

def test_katz_toy(cugraph):
    """test_katz_centrality.py:152-173."""
    G = _toy(cugraph)
    ck = cugraph.katz_centrality(G, alpha=0.01, beta=1.0, tol=0.000001, max_iter=1000)
    assert ck.columns == ["vertex", "katz_centrality"]
    centralities = [0.410614, 0.403211, 0.390689, 0.415175, 0.395125, 0.433226]
    assert np.abs(_scores(ck, "katz_centrality", 6) - centralities).max() <= 1e-2
    assert gv.nearly_equal(_scores(ck, "katz_centrality", 6), gv.G6_KATZ["result"])


def test_katz_default_alpha_and_argument_checks(cugraph):
    src, dst, w = _karate()
    G = cugraph.Graph()
    G.from_cudf_edgelist({"0": src, "1": dst}, "0", "1", store_transposed=True)
    degree_max = int(G.degree()["degree"].max())
    el = G.edgelist.edgelist_df
    want, _ = util.oracle_katz(el["src"].numpy(), el["dst"].numpy(), None, 34, dtype=np.float32,
                               alpha=1 / degree_max, beta=1.0, epsilon=1e-6, max_iterations=100,
                               normalize=True)
    got = _scores(cugraph.katz_centrality(G), "katz_centrality", 34)
    assert np.abs(got - want).max() <= 1e-6
    # nstart travels in the unused betas slot: accepted, no effect
    got2 = _scores(cugraph.katz_centrality(G, nstart={"vertex": np.arange(34),
                                                      "values": np.ones(34)}),
                   "katz_centrality", 34)
    assert np.array_equal(got, got2)
    with pytest.raises(ValueError, match="'alpha' must be a positive float or None"):
        cugraph.katz_centrality(G, alpha=-0.1)
    with pytest.raises(ValueError, match="'beta' must be a positive float"):
        cugraph.katz_centrality(G, alpha=0.1, beta=1)
    with pytest.raises(ValueError, match="'max_iter' must be a positive integer"):
        cugraph.katz_centrality(G, alpha=0.1, max_iter=0)
    with pytest.raises(ValueError, match="'tol' must be a positive float"):
        cugraph.katz_centrality(G, alpha=0.1, tol=0)


def test_eigenvector_toy(cugraph):
    """test_eigenvector_centrality.py:136-153 (graph = the undirected toy of the C-API test)."""
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist({"s": gv.E6_SRC, "d": gv.E6_DST, "w": gv.E6_WGT}, "s", "d", "w",
                         store_transposed=True)
    ck = cugraph.eigenvector_centrality(G, tol=1e-6, max_iter=200)
    assert ck.columns == ["vertex", "eigenvector_centrality"]
    centralities = [0.236325, 0.292055, 0.458457, 0.60533, 0.190498, 0.495942]
    assert np.abs(_scores(ck, "eigenvector_centrality", 6) - centralities).max() <= 1e-4
    with pytest.raises(ValueError, match="'max_iter' must be a positive integer"):
        cugraph.eigenvector_centrality(G, max_iter=-1)
    with pytest.raises(ValueError, match="'tol' must be a positive float"):
        cugraph.eigenvector_centrality(G, tol=1)


def test_hits_toy_and_nstart(cugraph):
    """c_api/hits_test.c golden vectors through the Python caller."""
    G = cugraph.Graph(directed=True)
    G.from_cudf_edgelist({"s": gv.G6_SRC, "d": gv.G6_DST, "w": gv.G6_WGT}, "s", "d", "w",
                         store_transposed=True)
    p = gv.G6_HITS
    df = cugraph.hits(G, max_iter=p["max_iterations"], tol=p["epsilon"], normalized=False)
    assert df.columns == ["vertex", "hubs", "authorities"]
    assert gv.nearly_equal(_scores(df, "hubs", 6), p["hubs"])
    assert gv.nearly_equal(_scores(df, "authorities", 6), p["authorities"])
    q = gv.G6_HITS_INITIAL
    nstart = {"vertex": q["initial_vertices"], "values": q["initial_hubs"]}  # float64 values
    df = cugraph.hits(G, max_iter=q["max_iterations"], tol=q["epsilon"], nstart=nstart,
                      normalized=False)
    assert gv.nearly_equal(_scores(df, "hubs", 6), q["hubs"])
    assert gv.nearly_equal(_scores(df, "authorities", 6), q["authorities"])
    # normalized=True (the Python default): scores sum to 1, same proportions
    h = _scores(cugraph.hits(G, max_iter=20, tol=p["epsilon"]), "hubs", 6)
    assert abs(h.sum() - 1.0) < 1e-6 and gv.nearly_equal(h / h.max(), p["hubs"])
