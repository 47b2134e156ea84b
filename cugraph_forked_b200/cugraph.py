"""cugraph-shaped Python callers of the pull path.

Same names, arguments, warnings and error behaviour as the reference's Python layer, with torch
tensors where the reference has cudf columns (there is no cudf or cupy dependency here):
  Graph / MultiGraph / DiGraph   python/cugraph/cugraph/structure/graph_classes.py:39,660,669
  Graph.from_cudf_edgelist       graph_classes.py:95 -> simpleGraph.py:110 (__from_edgelist)
  symmetrize / duplicate drop    structure/symmetrize.py:19-93 (concat + groupby.min)
  NumberMap                      structure/number_map.py (renumber, add_internal_vertex_id,
                                 unrenumber)
  _make_plc_graph                simpleGraph.py:806-868
  degree / in_degree / ...       simpleGraph.py:619-804, graph_primtypes_wrapper.pyx:106-248
  pagerank                       link_analysis/pagerank.py:61-244
  katz_centrality                centrality/katz_centrality.py
  eigenvector_centrality         centrality/eigenvector_centrality.py
  hits                           link_analysis/hits.py

Tables are `Frame` objects: an ordered set of equal-length columns, each a torch tensor on the
GPU (numeric data) or a host numpy array (strings and other non-numeric vertex labels, which
only ever pass through the NumberMap).  Anything with columns is accepted as input - a Frame, a
pandas.DataFrame, or a dict of arrays / tensors.

All graph preparation (symmetrize, duplicate drop, renumbering, degrees) is torch tensor code
that runs on the device the columns live on; every algorithm call goes to the CUDA library
through `pylibcugraph` (the cugraph_c C ABI).  There is no CPU implementation of the algorithms
in this module.
"""
from __future__ import annotations

import warnings

import numpy as np
import torch

from . import pylibcugraph as _plc

# Device the tables live on.  Always "cuda" in use; the CPU test-suite points it at "cpu"
# together with a stand-in for `_plc` to exercise the table logic without a GPU.
_DEVICE = "cuda"

_MAX_SG_EDGES = 2147483100  # simpleGraph.py:161

_handles = {}


def _resource_handle():
    """The reference makes a fresh `ResourceHandle()` for every graph and algorithm call; here one
    is made on first use and shared, so that a graph and the calls on it run on the same library
    stream and the library's per-handle workspace is kept between calls."""
    key = _plc.__name__
    if key not in _handles:
        _handles[key] = _plc.ResourceHandle()
    return _handles[key]


# --------------------------------------------------------------------------------------------
# Frame: the column table standing in for cudf.DataFrame
# --------------------------------------------------------------------------------------------

def _column(obj):
    """One column: numeric data becomes a torch tensor on `_DEVICE`, anything else a host numpy
    array.  Nulls are rejected as graph_classes.py:28 null_check does."""
    if isinstance(obj, torch.Tensor):
        if obj.dim() != 1:
            raise ValueError("columns must be one-dimensional")
        return obj.to(_DEVICE)
    if hasattr(obj, "__cuda_array_interface__"):
        return torch.as_tensor(obj, device="cuda").to(_DEVICE)
    if hasattr(obj, "isna") and bool(obj.isna().any()):
        raise ValueError("Series contains NULL values")
    if hasattr(obj, "to_numpy"):
        obj = obj.to_numpy()
    a = np.asarray(obj)
    if a.ndim != 1:
        raise ValueError("columns must be one-dimensional")
    if a.dtype.kind == "u":
        a = a.astype(np.int64 if a.dtype.itemsize >= 4 else np.int32)
    if a.dtype.kind in "ifb":
        a = np.ascontiguousarray(a)
        return torch.from_numpy(a if a.flags.writeable else a.copy()).to(_DEVICE)
    return a


def _take(col, index):
    if isinstance(col, torch.Tensor):
        return col[index.to(col.device)]
    return col[index.cpu().numpy()]


class Frame:
    """Ordered, equal-length columns.  The subset of the cudf.DataFrame surface the callers of
    this path use: column get / set, column lists, boolean-mask rows, rename, drop,
    sort_values, astype, to_pandas."""

    def __init__(self, data=None):
        self._cols = {}
        if data is None:
            return
        if isinstance(data, Frame):
            items = list(data._cols.items())
        elif isinstance(data, dict):
            items = list(data.items())
        elif hasattr(data, "columns"):
            items = [(c, data[c]) for c in data.columns]
        else:
            raise TypeError("input should be a Frame, a pandas.DataFrame or a dict of columns")
        for k, v in items:
            self[k] = v

    @property
    def columns(self):
        return list(self._cols)

    def __len__(self):
        for c in self._cols.values():
            return int(c.shape[0])
        return 0

    def __contains__(self, name):
        return name in self._cols

    def __iter__(self):
        return iter(self._cols)

    def items(self):
        return self._cols.items()

    def __setitem__(self, name, value):
        col = _column(value)
        for other, c in self._cols.items():
            if other != name and int(c.shape[0]) != int(col.shape[0]):
                raise ValueError(f"column '{name}' has {int(col.shape[0])} rows, table has "
                                 f"{int(c.shape[0])}")
        self._cols[name] = col

    def __getitem__(self, key):
        if isinstance(key, list):
            out = Frame()
            for k in key:
                out._cols[k] = self._cols[k]
            return out
        if isinstance(key, torch.Tensor) and key.dtype == torch.bool:
            return self.take(torch.nonzero(key, as_tuple=False).flatten())
        return self._cols[key]

    def take(self, index):
        out = Frame()
        for k, c in self._cols.items():
            out._cols[k] = _take(c, index)
        return out

    def rename(self, columns):
        out = Frame()
        for k, c in self._cols.items():
            out._cols[columns.get(k, k)] = c
        return out

    def drop(self, columns):
        columns = columns if isinstance(columns, list) else [columns]
        out = Frame()
        for k, c in self._cols.items():
            if k not in columns:
                out._cols[k] = c
        return out

    def astype(self, dtypes):
        out = Frame(self)
        for k, dt in dtypes.items():
            out._cols[k] = out._cols[k].to(_torch_dtype(dt))
        return out

    def sort_values(self, by):
        col = self._cols[by]
        if isinstance(col, torch.Tensor):
            order = torch.argsort(col, stable=True)
        else:
            order = torch.from_numpy(np.argsort(col, kind="stable"))
        return self.take(order)

    def reset_index(self, drop=True):
        return self

    def to_pandas(self):
        import pandas as pd
        return pd.DataFrame({k: (c.cpu().numpy() if isinstance(c, torch.Tensor) else c)
                             for k, c in self._cols.items()})

    def __repr__(self):
        return f"Frame({len(self)} rows: " + ", ".join(
            f"{k}:{c.dtype}" for k, c in self._cols.items()) + ")"


_NP_TO_TORCH = {"float32": torch.float32, "float64": torch.float64, "int32": torch.int32,
                "int64": torch.int64}


def _torch_dtype(dt):
    if isinstance(dt, torch.dtype):
        return dt
    return _NP_TO_TORCH[np.dtype(dt).name]


def _np_dtype(dt):
    return np.dtype(str(dt).replace("torch.", ""))


def _is_int_column(col):
    return isinstance(col, torch.Tensor) and col.dtype in (torch.int32, torch.int64)


def _as_list(x):
    return list(x) if isinstance(x, (list, tuple)) else [x]


# --------------------------------------------------------------------------------------------
# NumberMap: external vertex labels <-> contiguous internal ids
# --------------------------------------------------------------------------------------------

class NumberMap:
    """structure/number_map.py.  Internal id = rank of the label among the sorted distinct
    labels of the source and destination columns.  Single integer columns are handled with
    torch (sort / unique / searchsorted on the device); string, float and multi-column labels
    with a host pandas index, as they cannot live in a tensor."""

    def __init__(self, src_cols, dst_cols, col_names):
        self.col_names = [str(i) for i in range(len(col_names))]
        self.src_col_names = list(col_names)
        self.numbered = True
        self._tensor_keys = None
        self._index = None
        if len(src_cols) == 1 and _is_int_column(src_cols[0]) and _is_int_column(dst_cols[0]):
            both = torch.cat([src_cols[0].to(torch.int64), dst_cols[0].to(torch.int64)])
            self._tensor_keys = torch.unique(both, sorted=True).to(
                torch.promote_types(src_cols[0].dtype, dst_cols[0].dtype))
            n = int(self._tensor_keys.shape[0])
        else:
            import pandas as pd
            host = lambda cols: pd.DataFrame({i: (c.cpu().numpy() if isinstance(c, torch.Tensor)
                                                  else c) for i, c in enumerate(cols)})
            keys = pd.concat([host(src_cols), host(dst_cols)], ignore_index=True)
            keys = keys.drop_duplicates().sort_values(list(keys.columns)).reset_index(drop=True)
            self._keys_df = keys
            self._index = (pd.Index(keys[0]) if len(src_cols) == 1
                           else pd.MultiIndex.from_frame(keys))
            n = len(keys)
        self.id_type = torch.int32 if n <= np.iinfo(np.int32).max else torch.int64
        self.num_vertices = n

    def vertex_column_size(self):
        return len(self.col_names)

    def to_internal_vertex_id(self, cols):
        """Labels -> internal ids; -1 where a label is not a vertex of the graph."""
        cols = _as_list(cols)
        if len(cols) != len(self.col_names):
            raise ValueError(f"expected {len(self.col_names)} vertex column(s), got {len(cols)}")
        if self._tensor_keys is not None:
            q = cols[0]
            if not _is_int_column(q):
                raise ValueError("vertex ids must be integers for this graph")
            keys = self._tensor_keys.to(torch.int64)
            q = q.to(torch.int64)
            pos = torch.searchsorted(keys, q)
            pos_c = pos.clamp(max=max(self.num_vertices - 1, 0))
            hit = keys[pos_c] == q if self.num_vertices else torch.zeros_like(q, dtype=torch.bool)
            return torch.where(hit, pos_c, torch.full_like(pos_c, -1)).to(self.id_type)
        import pandas as pd
        host = [c.cpu().numpy() if isinstance(c, torch.Tensor) else np.asarray(c) for c in cols]
        target = pd.Index(host[0]) if len(host) == 1 else pd.MultiIndex.from_arrays(host)
        ids = self._index.get_indexer(target)
        return torch.from_numpy(np.ascontiguousarray(ids)).to(self.id_type).to(_DEVICE)

    def from_internal_vertex_id(self, ids):
        """Internal ids -> list of label columns (one per vertex column)."""
        if self._tensor_keys is not None:
            return [self._tensor_keys[ids.to(torch.int64)]]
        host = ids.cpu().numpy()
        out = []
        for i in range(len(self.col_names)):
            col = self._keys_df[i].to_numpy()[host]
            out.append(_column(col))
        return out

    def add_internal_vertex_id(self, df, id_column_name="id", col_names=None, drop=False):
        """number_map.py add_internal_vertex_id: `df` plus a column of internal ids looked up
        from the label column(s) `col_names`; with drop=True the label columns are removed and
        the id column comes first."""
        df = Frame(df)
        names = _as_list(col_names if col_names is not None else df.columns[:len(self.col_names)])
        ids = self.to_internal_vertex_id([df[c] for c in names])
        out = Frame()
        if drop:
            out._cols[id_column_name] = ids
            for k, c in df.items():
                if k not in names:
                    out._cols[k] = c
        else:
            for k, c in df.items():
                out._cols[k] = c
            out._cols[id_column_name] = ids
        return out

    def unrenumber(self, df, column_name):
        """number_map.py unrenumber: replace the internal ids in `column_name` with the original
        labels.  One label column keeps the name; several become "0_<name>", "1_<name>", ..."""
        df = Frame(df)
        labels = self.from_internal_vertex_id(df[column_name])
        out = Frame()
        if len(labels) == 1:
            for k, c in df.items():
                out._cols[k] = labels[0] if k == column_name else c
        else:
            for i, lab in enumerate(labels):
                out._cols[f"{i}_{column_name}"] = lab
            for k, c in df.items():
                if k != column_name:
                    out._cols[k] = c
        return out


# --------------------------------------------------------------------------------------------
# symmetrize + duplicate drop
# --------------------------------------------------------------------------------------------

def symmetrize(src, dst, values=None, multi=False, symmetrize=True):
    """structure/symmetrize.py:19-93 on tensors.  `values` is None or a list of attribute
    columns.  With symmetrize=True every edge is mirrored; unless `multi`, parallel edges are
    merged keeping the minimum of each attribute (groupby(...).min()), which also leaves the
    edge list ordered by (src, dst).  Returns (src, dst, values)."""
    values = None if values is None else list(values)
    if symmetrize:
        src, dst = torch.cat([src, dst]), torch.cat([dst, src])
        if values is not None:
            values = [torch.cat([v, v]) for v in values]
    if multi or src.numel() == 0:
        return src, dst, values
    order = torch.argsort(dst, stable=True)
    order = order[torch.argsort(src[order], stable=True)]
    s, d = src[order], dst[order]
    first = torch.ones_like(s, dtype=torch.bool)
    first[1:] = (s[1:] != s[:-1]) | (d[1:] != d[:-1])
    if values is not None:
        if bool(first.all()):
            values = [v[order] for v in values]
        else:
            group = torch.cumsum(first, 0) - 1
            n = int(group[-1]) + 1
            merged = []
            for v in values:
                vs = v[order]
                m = torch.empty(n, dtype=vs.dtype, device=vs.device)
                m.scatter_reduce_(0, group, vs, "amin", include_self=False)
                merged.append(m)
            values = merged
    return s[first], d[first], values


# --------------------------------------------------------------------------------------------
# Graph
# --------------------------------------------------------------------------------------------

class _EdgeList:
    """simpleGraph.py:39 EdgeList: `edgelist_df` with columns src, dst[, weights]."""

    def __init__(self, src, dst, weights=None):
        self.edgelist_df = Frame()
        self.edgelist_df._cols["src"] = src
        self.edgelist_df._cols["dst"] = dst
        self.weights = weights is not None
        if weights is not None:
            self.edgelist_df._cols["weights"] = weights


class Graph:
    """graph_classes.py:39.  `Graph()` is undirected (the edge list is symmetrized when it is
    loaded); `Graph(directed=True)` keeps the edges as given."""

    class Properties:
        def __init__(self, directed):
            self.directed = directed
            self.weights = False

    def __init__(self, m_graph=None, directed=False):
        self.graph_properties = Graph.Properties(directed)
        self.multi_edge = False
        self._reset()
        if m_graph is not None:
            if not isinstance(m_graph, MultiGraph):
                raise TypeError("m_graph can only be an instance of a cugraph.MultiGraph")
            elist = m_graph.view_edge_list()
            self.from_cudf_edgelist(elist, source="src", destination="dst",
                                    edge_attr="weights" if m_graph.is_weighted() else None)

    def _reset(self):
        self.edgelist = None
        self.renumber_map = None
        self.renumbered = False
        self.store_transposed = False
        self._plc_graph = None
        self._handle = None
        self.distributed = False
        self._node_count = None
        self._edge_count = None

    # ---- loading -------------------------------------------------------------------------

    def _edge_columns(self, input_df, source, destination, edge_attr):
        """The column checks of simpleGraph.py:120-160, same messages."""
        if not (isinstance(input_df, (Frame, dict)) or hasattr(input_df, "columns")):
            raise TypeError("input should be a Frame, a pandas.DataFrame or a dict of columns")
        available = set(input_df.keys() if isinstance(input_df, dict) else input_df.columns)
        s_col, d_col = _as_list(source), _as_list(destination)
        if not (set(s_col) <= available and set(d_col) <= available):
            raise ValueError("source column names and/or destination column names not found in "
                             "input. Recheck the source and destination parameters")
        if len(s_col) != len(d_col):
            raise ValueError("source and destination must name the same number of columns")
        attrs = None
        if edge_attr is not None:
            attrs = _as_list(edge_attr)
            if not set(attrs) <= available:
                raise ValueError("edge_attr column name not found in input."
                                 "Recheck the edge_attr parameter")
            if len(attrs) not in (1, 3):
                raise ValueError(f"Invalid number of edge attributes passed. {attrs}")
            if len(attrs) == 3 and not self.graph_properties.directed:
                raise ValueError("User-provided edge ids and edge types are not permitted for "
                                 "an undirected graph.")
        src_cols = [_column(input_df[c]) for c in s_col]
        dst_cols = [_column(input_df[c]) for c in d_col]
        values = None if attrs is None else [_column(input_df[c]) for c in attrs]
        return s_col, d_col, src_cols, dst_cols, values

    def from_cudf_edgelist(self, input_df, source="source", destination="destination",
                           edge_attr=None, renumber=True, store_transposed=False,
                           legacy_renum_only=False):
        """graph_classes.py:95 / simpleGraph.py:110.  `input_df`: Frame, pandas.DataFrame or
        dict of columns.  `source` / `destination` may be lists of column names (multi-column
        vertex labels, which need renumber=True)."""
        if self.edgelist is not None:
            raise RuntimeError("Graph already has values")
        s_col, d_col, src_cols, dst_cols, values = self._edge_columns(input_df, source,
                                                                      destination, edge_attr)
        if int(src_cols[0].shape[0]) > _MAX_SG_EDGES:
            raise ValueError("edge list is too big to fit in a single GPU")

        self.store_transposed = bool(store_transposed)
        plain_ints = (len(s_col) == 1 and _is_int_column(src_cols[0])
                      and _is_int_column(dst_cols[0]))
        if renumber:
            if plain_ints:
                # number_map.py renumber_and_segment 'skip_renumbering': integer labels are
                # renumbered inside the library and results come back with the labels
                src, dst = src_cols[0], dst_cols[0]
            else:
                self.renumber_map = NumberMap(src_cols, dst_cols, s_col)
                src = self.renumber_map.to_internal_vertex_id(src_cols)
                dst = self.renumber_map.to_internal_vertex_id(dst_cols)
                self.renumbered = True
        else:
            if len(s_col) > 1:
                raise ValueError("set renumber to True for multi column ids")
            if not plain_ints:
                raise ValueError("set renumber to True for non integer columns ids")
            src, dst = src_cols[0], dst_cols[0]
        if src.dtype != dst.dtype:
            common = torch.promote_types(src.dtype, dst.dtype)
            src, dst = src.to(common), dst.to(common)

        if values is not None:
            for v in values:
                if not isinstance(v, torch.Tensor):
                    raise ValueError("edge attributes must be numeric")
        src, dst, values = symmetrize(src, dst, values, multi=self.multi_edge,
                                      symmetrize=not self.graph_properties.directed)
        if values is not None:
            self.graph_properties.weights = True
        self.edgelist = _EdgeList(src, dst, None if values is None else values[0])
        try:
            self._make_plc_graph(values, store_transposed, renumber)
        except Exception:
            self._reset()  # a graph the library refused holds no values
            raise

    from_edgelist = from_cudf_edgelist

    def from_distributed_edgelist(self, input_df, source="source", destination="destination",
                                  edge_attr=None, resource_handle=None, store_transposed=False,
                                  num_global_edges=-1):
        """The multi-GPU loader: what Graph.from_dask_cudf_edgelist (graph_classes.py:228,
        simpleDistributedGraph.py) is to Dask, for one process per GPU.  Every rank calls it
        with ITS slice of the edge list (any split) and a `pylibcugraph.ResourceHandle` made
        from a multi-GPU engine handle; the library shuffles the edges to their 2D owners and
        renumbers (MGGraph, graphs.pyx:300).  Integer single-column ids; an undirected graph
        mirrors the local slice; parallel edges are kept as given (the reference's Dask
        groupby de-duplicates across workers - pass a de-duplicated list).  The centrality
        callers then return this rank's part of the result, and (vertex, value) inputs must
        name vertices whose result rows this rank holds."""
        if self.edgelist is not None:
            raise RuntimeError("Graph already has values")
        if resource_handle is None:
            raise ValueError("from_distributed_edgelist needs the multi-GPU resource_handle")
        s_col, d_col, src_cols, dst_cols, values = self._edge_columns(input_df, source,
                                                                      destination, edge_attr)
        if len(s_col) > 1 or not (_is_int_column(src_cols[0]) and _is_int_column(dst_cols[0])):
            raise ValueError("the distributed loader takes single integer id columns")
        if values is not None and len(values) != 1:
            raise ValueError("the distributed loader takes one edge attribute (the weight)")
        src, dst = src_cols[0], dst_cols[0]
        if src.dtype != dst.dtype:
            common = torch.promote_types(src.dtype, dst.dtype)
            src, dst = src.to(common), dst.to(common)
        weight = None if values is None else values[0]
        if weight is not None and weight.dtype in (torch.int32, torch.int64):
            weight = weight.to(torch.float32 if weight.dtype == torch.int32 else torch.float64)
        if not self.graph_properties.directed:
            off_diagonal = src != dst
            src, dst = (torch.cat([src, dst[off_diagonal]]), torch.cat([dst, src[off_diagonal]]))
            if weight is not None:
                weight = torch.cat([weight, weight[off_diagonal]])
        self.store_transposed = bool(store_transposed)
        self.graph_properties.weights = weight is not None
        self.edgelist = _EdgeList(src, dst, weight)
        self.distributed = True
        self._handle = resource_handle
        props = _plc.GraphProperties(is_multigraph=True,
                                     is_symmetric=not self.graph_properties.directed)
        try:
            self._plc_graph = _plc.MGGraph(
                resource_handle=resource_handle, graph_properties=props,
                src_array=src.contiguous(), dst_array=dst.contiguous(),
                weight_array=None if weight is None else weight.contiguous(),
                store_transposed=store_transposed, num_edges=num_global_edges,
                do_expensive_check=False)
        except Exception:
            self._reset()
            raise

    def from_cudf_adjlist(self, offset_col, index_col, value_col=None, renumber=True,
                          store_transposed=False):
        """graph_classes.py:173: compressed-sparse-row input - `offset_col` [V + 1],
        `index_col` [E] destinations, optional `value_col` [E].  The reference only stores the
        arrays here (simpleGraph.py:365-370, a graph loaded this way cannot be handed to its
        pylibcugraph algorithms); here the rows are expanded on the device and the graph is
        built as from an edge list, so the centrality callers work on it."""
        if self.edgelist is not None:
            raise RuntimeError("Graph already has values")
        offsets, indices = _column(offset_col), _column(index_col)
        if not (_is_int_column(offsets) and _is_int_column(indices)):
            raise ValueError("offset_col and index_col must be integer columns")
        if offsets.numel() == 0 or int(offsets[-1]) != indices.numel() or int(offsets[0]) != 0:
            raise ValueError("offset_col must run from 0 to len(index_col)")
        degrees = offsets[1:] - offsets[:-1]
        if bool((degrees < 0).any()):
            raise ValueError("offset_col must be non-decreasing")
        src = torch.repeat_interleave(
            torch.arange(degrees.numel(), dtype=indices.dtype, device=indices.device),
            degrees.to(torch.int64))
        df = {"src": src, "dst": indices}
        if value_col is not None:
            df["weights"] = _column(value_col)
        self.from_cudf_edgelist(df, "src", "dst", "weights" if value_col is not None else None,
                                renumber=renumber, store_transposed=store_transposed)

    def _compressed(self, by_destination):
        df = self._need_edges()
        major, minor = (df["dst"], df["src"]) if by_destination else (df["src"], df["dst"])
        n = self.number_of_vertices()
        order = torch.argsort(minor, stable=True)
        order = order[torch.argsort(major[order], stable=True)]
        counts = torch.bincount(major.to(torch.int64), minlength=n)
        offsets = torch.zeros(n + 1, dtype=torch.int64, device=major.device)
        offsets[1:] = torch.cumsum(counts, 0)
        weights = df["weights"][order] if self.is_weighted() else None
        return offsets.to(major.dtype), minor[order], weights

    def view_adj_list(self):
        """simpleGraph.py:372-414: (offsets, indices, weights) of the stored edge list by
        source - internal ids, neighbours ascending; weights None for an unweighted graph."""
        return self._compressed(False)

    def view_transposed_adj_list(self):
        """simpleGraph.py:416-465: the same by destination (the pull direction)."""
        return self._compressed(True)

    def from_pandas_edgelist(self, pdf, source="source", destination="destination",
                             edge_attr=None, renumber=True, store_transposed=False):
        """graph_classes.py:295."""
        import pandas as pd
        if not isinstance(pdf, pd.core.frame.DataFrame):
            raise TypeError("pdf input is not a Pandas DataFrame")
        self.from_cudf_edgelist(pdf, source=source, destination=destination, edge_attr=edge_attr,
                                renumber=renumber, store_transposed=store_transposed)

    def from_numpy_array(self, np_array, nodes=None):
        """graph_classes.py:368: dense adjacency matrix; non-zero entries are the edges."""
        if not isinstance(np_array, np.ndarray):
            raise TypeError("np_array input is not a Numpy array")
        if len(np_array.shape) != 2:
            raise ValueError("np_array is not a 2D matrix")
        src, dst = np_array.nonzero()
        weight = np_array[src, dst]
        df = {"src": src if nodes is None else np.asarray(nodes)[src],
              "dst": dst if nodes is None else np.asarray(nodes)[dst], "weight": weight}
        self.from_cudf_edgelist(df, "src", "dst", edge_attr="weight")

    def from_pandas_adjacency(self, pdf):
        """graph_classes.py:352."""
        import pandas as pd
        if not isinstance(pdf, pd.core.frame.DataFrame):
            raise TypeError("pdf input is not a Pandas DataFrame")
        self.from_numpy_array(pdf.to_numpy(), pdf.columns)

    def _make_plc_graph(self, value_col=None, store_transposed=False, renumber=True):
        """simpleGraph.py:806-868.  Integer weights are widened to the float type of the same
        size.  Where the reference hands the library a column of float32 ones for an unweighted
        edge list, no weight array is passed: the library's unweighted graph computes in float32
        with implicit unit weights, so results are the same and the edge values are not read."""
        weight_col = id_col = type_col = None
        if value_col is not None:
            if len(value_col) == 3:
                weight_col, id_col, type_col = value_col
            else:
                weight_col = value_col[0]
            if weight_col.dtype == torch.int32:
                weight_col = weight_col.to(torch.float32)
            elif weight_col.dtype == torch.int64:
                weight_col = weight_col.to(torch.float64)
        props = _plc.GraphProperties(is_multigraph=self.multi_edge,
                                     is_symmetric=not self.graph_properties.directed)
        df = self.edgelist.edgelist_df
        self._handle = _resource_handle()
        self._plc_graph = _plc.SGGraph(
            resource_handle=self._handle, graph_properties=props,
            src_array=df["src"].contiguous(), dst_array=df["dst"].contiguous(),
            weight_array=None if weight_col is None else weight_col.contiguous(),
            edge_id_array=id_col, edge_type_array=type_col, store_transposed=store_transposed,
            renumber=renumber, do_expensive_check=False)

    def clear(self):
        """graph_classes.py:523."""
        self._reset()

    # ---- properties ----------------------------------------------------------------------

    def is_directed(self):
        return self.graph_properties.directed

    def is_weighted(self):
        return self.graph_properties.weights

    def is_renumbered(self):
        return self.renumbered

    def is_multigraph(self):
        return self.multi_edge

    def _need_edges(self, local_ok=False):
        if self.edgelist is None:
            raise RuntimeError("Graph is Empty")
        if self.distributed and not local_ok:
            raise NotImplementedError("this rank holds only its slice of a distributed graph")
        return self.edgelist.edgelist_df

    def number_of_vertices(self):
        """simpleGraph.py:568: largest id in the edge list plus one."""
        if self._node_count is None:
            df = self._need_edges()
            if len(df) == 0:
                self._node_count = 0
            else:
                self._node_count = int(max(int(df["src"].max()), int(df["dst"].max())) + 1)
        return self._node_count

    number_of_nodes = number_of_vertices

    def number_of_edges(self, directed_edges=False):
        """simpleGraph.py:592: an undirected graph counts each mirrored pair once."""
        df = self._need_edges()
        if directed_edges:
            return len(df)
        if self._edge_count is None:
            if self.graph_properties.directed:
                self._edge_count = len(df)
            else:
                self._edge_count = int((df["src"] >= df["dst"]).sum())
        return self._edge_count

    def view_edge_list(self):
        """simpleGraph.py:310: the stored edge list with original labels; an undirected graph
        shows one direction of each pair (src <= dst on internal ids)."""
        df = self._need_edges()
        if not self.graph_properties.directed:
            df = df[df["src"] <= df["dst"]]
        if self.renumbered:
            df = self.renumber_map.unrenumber(df, "src")
            df = self.renumber_map.unrenumber(df, "dst")
        return df

    def edges(self):
        return self.view_edge_list().drop("weights") if self.is_weighted() \
            else self.view_edge_list()

    def nodes(self):
        """simpleGraph.py:977: the distinct vertex labels."""
        df = self._need_edges()
        if self.renumbered:
            ids = torch.arange(self.renumber_map.num_vertices, device=df["src"].device)
            labels = self.renumber_map.from_internal_vertex_id(ids)
            return labels[0] if len(labels) == 1 else labels
        return torch.unique(torch.cat([df["src"], df["dst"]]))

    def has_node(self, n):
        """simpleGraph.py:928."""
        df = self._need_edges()
        if self.renumbered:
            q = self.renumber_map.to_internal_vertex_id([_column(np.asarray(c)[None])
                                                         for c in _as_list(n)])
            return bool(q[0] >= 0)
        return bool(((df["src"] == n) | (df["dst"] == n)).any())

    def has_edge(self, u, v):
        """simpleGraph.py:939."""
        df = self._need_edges()
        if self.renumbered:
            lookup = lambda x: int(self.renumber_map.to_internal_vertex_id(
                [_column(np.asarray(c)[None]) for c in _as_list(x)])[0])
            u, v = lookup(u), lookup(v)
            if u < 0 or v < 0:
                return False
        return bool(((df["src"] == u) & (df["dst"] == v)).any())

    # ---- renumbering ---------------------------------------------------------------------

    def unrenumber(self, df, column_name, preserve_order=False, get_column_names=False):
        """graph_classes.py:412."""
        if self.renumber_map is None:
            return Frame(df)
        return self.renumber_map.unrenumber(df, column_name)

    def lookup_internal_vertex_id(self, df, column_name=None):
        """graph_classes.py:452: internal ids of the labels in `df` (a column, or a table with
        `column_name` naming the label column(s)); -1 for labels that are not vertices."""
        if self.renumber_map is None:
            raise ValueError("graph is not renumbered")
        if column_name is None:
            cols = [_column(df)] if not hasattr(df, "columns") else [
                _column(df[c]) for c in df.columns]
        else:
            cols = [_column(df[c]) for c in _as_list(column_name)]
        return self.renumber_map.to_internal_vertex_id(cols)

    def add_internal_vertex_id(self, df, internal_column_name, external_column_name, drop=True,
                               preserve_order=False):
        """graph_classes.py:476."""
        if self.renumber_map is None:
            raise ValueError("graph is not renumbered")
        return self.renumber_map.add_internal_vertex_id(df, internal_column_name,
                                                        external_column_name, drop)

    # ---- degrees -------------------------------------------------------------------------

    def _degree_columns(self):
        df = self._need_edges()
        n = self.number_of_vertices()
        dev = df["src"].device
        out_deg = torch.bincount(df["src"].to(torch.int64), minlength=n).to(torch.int32)
        in_deg = torch.bincount(df["dst"].to(torch.int64), minlength=n).to(torch.int32)
        return torch.arange(n, dtype=torch.int32, device=dev), in_deg, out_deg

    def _degree_frame(self, columns, vertex_subset):
        df = Frame()
        for k, c in columns.items():
            df._cols[k] = c
        if self.renumbered:
            df = self.renumber_map.unrenumber(df, "vertex")
        if vertex_subset is not None:
            sub = _column(np.asarray(vertex_subset) if not isinstance(
                vertex_subset, torch.Tensor) else vertex_subset)
            df = df[torch.isin(df["vertex"], sub.to(df["vertex"].dtype))]
        return df

    def in_degree(self, vertex_subset=None):
        """simpleGraph.py:619."""
        v, i, _ = self._degree_columns()
        return self._degree_frame({"vertex": v, "degree": i}, vertex_subset)

    def out_degree(self, vertex_subset=None):
        """simpleGraph.py:657."""
        v, _, o = self._degree_columns()
        return self._degree_frame({"vertex": v, "degree": o}, vertex_subset)

    def degree(self, vertex_subset=None):
        """simpleGraph.py:695: in-degree plus out-degree over the stored edge list, which for an
        undirected graph holds both directions (graph_primtypes_wrapper.pyx:115-116)."""
        v, i, o = self._degree_columns()
        return self._degree_frame({"vertex": v, "degree": i + o}, vertex_subset)

    def degrees(self, vertex_subset=None):
        """simpleGraph.py:735."""
        v, i, o = self._degree_columns()
        return self._degree_frame({"vertex": v, "in_degree": i, "out_degree": o}, vertex_subset)

    # ---- conversions ---------------------------------------------------------------------

    def _copy_as(self, G, directed, store_transposed):
        df = self._need_edges()
        G._reset()
        G.graph_properties = Graph.Properties(directed)
        G.graph_properties.weights = self.graph_properties.weights
        G.renumbered = self.renumbered
        G.renumber_map = self.renumber_map
        G.store_transposed = bool(store_transposed)
        values = [df["weights"]] if self.is_weighted() else None
        src, dst = df["src"], df["dst"]
        if not directed and self.graph_properties.directed:
            src, dst, values = symmetrize(src, dst, values, multi=G.multi_edge, symmetrize=True)
        G.edgelist = _EdgeList(src, dst, None if values is None else values[0])
        G._make_plc_graph(values, store_transposed, renumber=True)
        return G

    def to_directed(self, store_transposed=False):
        """graph_classes.py:585: a directed graph over the stored edges (both directions of
        every edge when this graph is undirected)."""
        G = MultiGraph(directed=True) if self.multi_edge else Graph(directed=True)
        return self._copy_as(G, True, store_transposed)

    def to_undirected(self, store_transposed=False):
        """graph_classes.py:614."""
        G = MultiGraph(directed=False) if self.multi_edge else Graph(directed=False)
        return self._copy_as(G, False, store_transposed)


class DiGraph(Graph):
    """graph_classes.py:660 (deprecated in the reference in favour of Graph(directed=True))."""

    def __init__(self, m_graph=None):
        warnings.warn("DiGraph is deprecated, use Graph(directed=True) instead",
                      DeprecationWarning)
        super().__init__(m_graph, directed=True)


class MultiGraph(Graph):
    """graph_classes.py:669: parallel edges are kept."""

    def __init__(self, directed=False):
        super().__init__(directed=directed)
        self.multi_edge = True


class MultiDiGraph(MultiGraph):
    """graph_classes.py:685."""

    def __init__(self):
        warnings.warn("MultiDiGraph is deprecated, use MultiGraph(directed=True) instead",
                      DeprecationWarning)
        super().__init__(directed=True)


def is_directed(G):
    return G.is_directed()


def is_weighted(G):
    return G.is_weighted()


def is_multigraph(G):
    return G.is_multigraph()


# --------------------------------------------------------------------------------------------
# algorithm callers
# --------------------------------------------------------------------------------------------

def _edge_attr_dtype(G):
    """pagerank.py:42-46: float32 for an unweighted graph, else the stored weight type (after
    the int -> float widening of _make_plc_graph)."""
    if not G.edgelist.weights:
        return torch.float32
    dt = G.edgelist.edgelist_df["weights"].dtype
    return {torch.int32: torch.float32, torch.int64: torch.float64}.get(dt, dt)


def renumber_vertices(input_graph, input_df):
    """pagerank.py:27-36: replace the label column(s) of a (vertex..., value) table with the
    internal ids, in a column called 'vertex'.  With multi-column labels every column but the
    last is a label column."""
    input_df = Frame(input_df)
    if input_graph.renumber_map.vertex_column_size() > 1:
        cols = input_df.columns[:-1]
    else:
        cols = "vertex"
    out = input_graph.add_internal_vertex_id(input_df, "vertex", cols)
    if bool((out["vertex"] < 0).any()):
        raise ValueError("a vertex of the input table is not in the graph")
    return out


def ensure_valid_dtype(input_graph, input_df, input_df_name):
    """pagerank.py:41-58: warn and cast when the 'values' column does not have the graph's
    edge_attr type."""
    edge_attr_dtype = _edge_attr_dtype(input_graph)
    input_df = Frame(input_df)
    input_df_dtype = input_df["values"].dtype
    if input_df_dtype != edge_attr_dtype:
        warning_msg = (f"PageRank requires '{input_df_name}' values "
                       "to match the graph's 'edge_attr' type. "
                       f"edge_attr type is: {_np_dtype(edge_attr_dtype)} and got "
                       f"'{input_df_name}' values of type: "
                       f"{_np_dtype(input_df_dtype)}.")
        warnings.warn(warning_msg, UserWarning)
        input_df = input_df.astype({"values": edge_attr_dtype})
    return input_df


def _nx():
    try:
        import networkx
    except ImportError:  # NetworkX is optional, as in the reference (utils.import_optional)
        return None
    return networkx


def convert_from_nx(nxG, weight=None, do_renumber=True):
    """utilities/nx_factory.py:76-106: NetworkX Graph / DiGraph -> Graph.  An unweighted graph
    gives columns src, dst; a weighted one takes the edge attribute `weight`, or, when `weight`
    is None, the single edge attribute the graph has (nx.to_pandas_edgelist), as float32
    (:55).  The graph is stored transposed - the reference stores it untransposed here and then
    warns about it in every centrality call on a NetworkX input."""
    nx = _nx()
    if isinstance(nxG, nx.classes.digraph.DiGraph):
        G = Graph(directed=True)
    elif isinstance(nxG, nx.classes.graph.Graph):
        G = Graph()
    else:
        raise TypeError(f"nxG must be either a NetworkX Graph or DiGraph, got {type(nxG)}")
    if nx.is_weighted(nxG) is False:
        edges = list(nxG.edges(data=False))
        df = {"src": _label_column([s for s, _ in edges]), "dst": _label_column([d for _, d in edges])}
        G.from_cudf_edgelist(df, source="src", destination="dst", edge_attr=None,
                             renumber=do_renumber, store_transposed=True)
    elif weight is None:
        pdf = nx.to_pandas_edgelist(nxG)
        wt_col = [c for c in pdf.columns if c not in ("source", "target")]
        if len(wt_col) != 1:
            raise ValueError("Unable to determine weight column name")
        pdf = pdf.rename(columns={wt_col[0]: "weight"})
        G.from_cudf_edgelist(pdf, source="source", destination="target", edge_attr="weight",
                             renumber=do_renumber, store_transposed=True)
    else:
        edges = list(nxG.edges(data=weight))
        df = {"src": _label_column([s for s, _, _ in edges]),
              "dst": _label_column([d for _, d, _ in edges]),
              "weight": np.asarray([w for _, _, w in edges], dtype=np.float32)}
        G.from_cudf_edgelist(df, source="src", destination="dst", edge_attr="weight",
                             renumber=do_renumber, store_transposed=True)
    return G


def _label_column(values):
    a = np.asarray(values)
    return a.astype(np.int64) if a.dtype.kind in "iu" else a


def ensure_cugraph_obj_for_nx(obj, nx_weight_attr="weight"):
    """utilities/utils.py:326-341: (Graph, is_networkx_input)."""
    nx = _nx()
    if nx is not None and isinstance(obj, nx.Graph):
        return convert_from_nx(obj, weight=nx_weight_attr), True
    if isinstance(obj, Graph):
        return obj, False
    raise TypeError(f"input must be either a cuGraph or NetworkX graph type, got {type(obj)}")


def df_score_to_dictionary(df, k, v="vertex"):
    """utilities/nx_factory.py:109-137: {vertex: score}, ordered by vertex."""
    df = df.sort_values(v)
    keys = df[v].tolist()
    return dict(zip(keys, df[k].tolist()))


def _check_graph(G, name):
    if G._plc_graph is None:
        raise RuntimeError("Graph is Empty")


def _warn_not_transposed(G, algo):
    if G.store_transposed is False:
        warnings.warn(f"{algo} expects the 'store_transposed' flag "
                      "to be set to 'True' for optimal performance during "
                      "the graph creation", UserWarning)


def _vertex_ids(G, col):
    """Vertex id column of an input table, ready for the library: contiguous and in the graph's
    vertex type (the C API rejects an id array whose type differs from the graph's)."""
    if not _is_int_column(col):
        raise ValueError("vertex ids must be integers")
    return col.to(G.edgelist.edgelist_df["src"].dtype).contiguous()


def pagerank(G, alpha=0.85, personalization=None, precomputed_vertex_out_weight=None,
             max_iter=100, tol=1.0e-5, nstart=None, weight=None, dangling=None):
    """link_analysis/pagerank.py:61-244.  Returns a Frame with columns 'vertex' and 'pagerank'
    (a {vertex: score} dict when G is a NetworkX graph).  `personalization` / `nstart`: tables
    with 'vertex' and 'values'; `precomputed_vertex_out_weight`: 'vertex' and 'sums'.  `weight`
    names the edge attribute of a NetworkX input; `dangling` is ignored, as in the reference."""
    G, isNx = ensure_cugraph_obj_for_nx(G, weight)
    _check_graph(G, "pagerank")
    _warn_not_transposed(G, "Pagerank")
    initial_guess_vertices = initial_guess_values = None
    pre_vtx_o_wgt_vertices = pre_vtx_o_wgt_sums = None
    do_expensive_check = False

    if nstart is not None:
        nstart = Frame(nstart)
        if G.renumbered is True:
            nstart = renumber_vertices(G, nstart)
        nstart = ensure_valid_dtype(G, nstart, "nstart")
        initial_guess_vertices = _vertex_ids(G, nstart["vertex"])
        initial_guess_values = nstart["values"].contiguous()

    if precomputed_vertex_out_weight is not None:
        precomputed_vertex_out_weight = Frame(precomputed_vertex_out_weight)
        if G.renumbered is True:
            precomputed_vertex_out_weight = renumber_vertices(G, precomputed_vertex_out_weight)
        pre_vtx_o_wgt_vertices = _vertex_ids(G, precomputed_vertex_out_weight["vertex"])
        pre_vtx_o_wgt_sums = precomputed_vertex_out_weight["sums"].to(
            _edge_attr_dtype(G)).contiguous()

    if personalization is not None:
        if not (isinstance(personalization, (Frame, dict)) or hasattr(personalization, "columns")):
            raise NotImplementedError("personalization other than a table of 'vertex' and "
                                      "'values' columns currently not supported")
        personalization = Frame(personalization)
        if G.renumbered is True:
            personalization = renumber_vertices(G, personalization)
        personalization = ensure_valid_dtype(G, personalization, "personalization")
        vertex, pagerank_values = _plc.personalized_pagerank(
            resource_handle=G._handle, graph=G._plc_graph,
            precomputed_vertex_out_weight_vertices=pre_vtx_o_wgt_vertices,
            precomputed_vertex_out_weight_sums=pre_vtx_o_wgt_sums,
            personalization_vertices=_vertex_ids(G, personalization["vertex"]),
            personalization_values=personalization["values"].contiguous(),
            initial_guess_vertices=initial_guess_vertices,
            initial_guess_values=initial_guess_values,
            alpha=alpha, epsilon=tol, max_iterations=max_iter,
            do_expensive_check=do_expensive_check)
    else:
        vertex, pagerank_values = _plc.pagerank(
            resource_handle=G._handle, graph=G._plc_graph,
            precomputed_vertex_out_weight_vertices=pre_vtx_o_wgt_vertices,
            precomputed_vertex_out_weight_sums=pre_vtx_o_wgt_sums,
            initial_guess_vertices=initial_guess_vertices,
            initial_guess_values=initial_guess_values,
            alpha=alpha, epsilon=tol, max_iterations=max_iter,
            do_expensive_check=do_expensive_check)

    df = Frame()
    df["vertex"] = vertex
    df["pagerank"] = pagerank_values
    if G.renumbered:
        df = G.unrenumber(df, "vertex")
    if isNx is True:
        df = df_score_to_dictionary(df, "pagerank")
    return df


def katz_centrality(G, alpha=None, beta=1.0, max_iter=100, tol=1.0e-6, nstart=None,
                    normalized=True):
    """centrality/katz_centrality.py.  Returns a Frame with 'vertex' and 'katz_centrality'.
    alpha=None uses 1 / (largest degree).  As in the reference, `normalized` is accepted and
    not forwarded, and `nstart` travels in the C API's `betas` slot, which the library accepts
    and does not use (c_api/katz.cpp:158)."""
    G, isNx = ensure_cugraph_obj_for_nx(G)
    _check_graph(G, "katz_centrality")
    _warn_not_transposed(G, "Katz centrality")

    if alpha is None:
        degree_max = int(G.degree()["degree"].max())
        alpha = 1 / degree_max

    if (alpha is not None) and (alpha <= 0.0):
        raise ValueError(f"'alpha' must be a positive float or None, got: {alpha}")
    elif (not isinstance(beta, float)) or (beta <= 0.0):
        raise ValueError(f"'beta' must be a positive float or None, got: {beta}")
    if (not isinstance(max_iter, int)) or (max_iter <= 0):
        raise ValueError(f"'max_iter' must be a positive integer, got: {max_iter}")
    if (not isinstance(tol, float)) or (tol <= 0.0):
        raise ValueError(f"'tol' must be a positive float, got: {tol}")

    betas = None
    if nstart is not None:
        nstart = Frame(nstart)
        if G.renumbered is True:
            if G.renumber_map.vertex_column_size() > 1:
                cols = nstart.columns[:-1]
            else:
                cols = "vertex"
            nstart = G.add_internal_vertex_id(nstart, "vertex", cols)
        betas = nstart[nstart.columns[0]].contiguous()

    vertices, values = _plc.katz_centrality(
        resource_handle=G._handle, graph=G._plc_graph, betas=betas, alpha=alpha,
        beta=beta, epsilon=tol, max_iterations=max_iter, do_expensive_check=False)

    df = Frame()
    df["vertex"] = vertices
    df["katz_centrality"] = values
    if G.renumbered:
        df = G.unrenumber(df, "vertex")
    if isNx is True:
        return df_score_to_dictionary(df, "katz_centrality")
    return df


def eigenvector_centrality(G, max_iter=100, tol=1.0e-6):
    """centrality/eigenvector_centrality.py.  Returns a Frame with 'vertex' and
    'eigenvector_centrality'."""
    if (not isinstance(max_iter, int)) or max_iter <= 0:
        raise ValueError(f"'max_iter' must be a positive integer, got: {max_iter}")
    if (not isinstance(tol, float)) or (tol <= 0.0):
        raise ValueError(f"'tol' must be a positive float, got: {tol}")
    G, isNx = ensure_cugraph_obj_for_nx(G)
    _check_graph(G, "eigenvector_centrality")
    _warn_not_transposed(G, "Eigenvector centrality")

    vertices, values = _plc.eigenvector_centrality(
        resource_handle=G._handle, graph=G._plc_graph, epsilon=tol,
        max_iterations=max_iter, do_expensive_check=False)

    df = Frame()
    df["vertex"] = vertices
    df["eigenvector_centrality"] = values
    if G.renumbered:
        df = G.unrenumber(df, "vertex")
    if isNx is True:
        return df_score_to_dictionary(df, "eigenvector_centrality")
    return df


def hits(G, max_iter=100, tol=1.0e-5, nstart=None, normalized=True):
    """link_analysis/hits.py.  Returns a Frame with 'vertex', 'hubs' and 'authorities'.
    `nstart`: table with 'vertex' and 'values', the starting hub scores.  The reference hands
    nstart['vertex'] to the library as is; with a label-renumbered graph the labels are
    translated first, which the reference leaves undone (and would fail on)."""
    G, isNx = ensure_cugraph_obj_for_nx(G)
    _check_graph(G, "hits")
    _warn_not_transposed(G, "HITS")
    init_hubs_guess_vertices = init_hubs_guess_values = None
    if nstart is not None:
        nstart = Frame(nstart)
        if G.renumbered is True:
            nstart = renumber_vertices(G, nstart)
        init_hubs_guess_vertices = _vertex_ids(G, nstart["vertex"])
        init_hubs_guess_values = nstart["values"].to(_edge_attr_dtype(G)).contiguous()

    vertices, hubs, authorities = _plc.hits(
        resource_handle=G._handle, graph=G._plc_graph, tol=tol, max_iter=max_iter,
        initial_hubs_guess_vertices=init_hubs_guess_vertices,
        initial_hubs_guess_values=init_hubs_guess_values, normalized=normalized,
        do_expensive_check=False)

    results = Frame()
    results["vertex"] = vertices
    results["hubs"] = hubs
    results["authorities"] = authorities
    if G.renumbered:
        results = G.unrenumber(results, "vertex")
    if isNx is True:
        results = (df_score_to_dictionary(results, "hubs"),
                   df_score_to_dictionary(results, "authorities"))
    return results
