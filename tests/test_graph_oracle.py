"""CPU: the numpy restatement of the integer graph-build path (oracle/graph_oracle.py)."""
import numpy as np
import pytest

from oracle import graph_oracle as go
from tests import util


def test_murmur3_known_answers():
    """MurmurHash3_x86_32 test vectors (seed 0) - pins the restatement of cuco's hash, which is
    not in the reference tree (parity unpinned against cuco itself)."""
    keys = np.array([0x00000000, 0x87654321 - (1 << 32), -1], dtype=np.int32)
    h = go.murmur3_32(keys)
    assert [int(x) for x in h] == [0x2362F9DE, 0xF55B516B, 0x76293B50]
    # 8-byte keys hash their 8 little-endian bytes: the zero-extended key is NOT the 4-byte hash
    h64 = go.murmur3_32(np.array([0, 0x87654321], dtype=np.int64))
    assert int(h64[0]) == 0x63852AFC and int(h64[1]) != 0xF55B516B


def test_grid_2d_matches_reference_choice():
    # P=2 -> 1x2, 4 -> 2x2, 8 -> 2x4 (row_comm_size x col_comm_size)
    assert [go.grid_2d(p) for p in (1, 2, 4, 8, 16)] == [(1, 1), (1, 2), (2, 2), (2, 4), (4, 4)]


def test_edge_owner_formulas():
    rng = np.random.default_rng(0)
    maj = rng.integers(0, 1 << 20, 1000).astype(np.int32)
    mnr = rng.integers(0, 1 << 20, 1000).astype(np.int32)
    P, (R, C) = 8, go.grid_2d(8)
    gpu = go.gpu_id_from_ext_edge_endpoints(maj, mnr, P, R, C)
    part = go.partition_id_from_ext_edge_endpoints(maj, mnr, P, R, C)
    mj, mn = go.gpu_id_from_ext_vertex(maj, P), go.gpu_id_from_ext_vertex(mnr, P)
    assert np.all((gpu >= 0) & (gpu < P))
    # the owner GPU sits in the column block of the minor's owner and the row of the major's
    assert np.array_equal(gpu // R, mn // R) and np.array_equal(gpu % R, mj % R)
    # local partition index on that GPU = major owner's column-block index
    assert np.array_equal(part % C, mn // R) and np.array_equal(part // C, mj)


def test_renumber_is_degree_descending_permutation():
    src, dst = util.rmat(10)
    verts = np.arange(1 << 10, dtype=np.int32)
    labels, seg, nmaj, nmin = go.renumber_sg(dst, src, verts)
    assert np.array_equal(np.sort(labels), verts)
    deg = np.bincount(nmaj, minlength=len(labels))
    assert np.all(np.diff(deg) <= 0)
    # ties keep ascending external id
    for d in np.unique(deg):
        ids = labels[deg == d]
        assert np.all(np.diff(ids) > 0)
    assert seg[0] == 0 and seg[-1] == len(labels)
    assert seg[1] == np.count_nonzero(deg >= 1024) and seg[2] == np.count_nonzero(deg >= 32)
    assert seg[3] == np.count_nonzero(deg >= 1)
    # round trip
    assert np.array_equal(labels[nmaj], dst) and np.array_equal(labels[nmin], src)


def test_compress_sorted():
    src, dst = util.rmat(8, edge_factor=8)
    labels, seg, nmaj, nmin = go.renumber_sg(dst, src)
    w = np.arange(len(src), dtype=np.float32)
    offs, idx, ww = go.compress_sorted(nmaj, nmin, w, 0, len(labels))
    assert offs[-1] == len(src)
    for r in range(len(labels)):
        row = idx[offs[r]:offs[r + 1]]
        assert np.all(np.diff(row) >= 0)
    # every (major, minor, weight) triple survives
    maj = np.repeat(np.arange(len(labels)), np.diff(offs))
    a = np.sort(maj.astype(np.int64) << 40 | idx.astype(np.int64) << 20 | ww.astype(np.int64))
    b = np.sort(nmaj.astype(np.int64) << 40 | nmin.astype(np.int64) << 20 | w.astype(np.int64))
    assert np.array_equal(a, b)


def test_rmat_quadrant_ratios():
    """as cpp/tests/generators/generate_rmat_test.cpp:33-140: top-level quadrant shares ~ a,b,c,d"""
    scale = 12
    src, dst = go.rmat_edgelist(scale, 1 << 18, a=0.57, b=0.19, c=0.19, seed=3)
    assert src.min() >= 0 and src.max() < (1 << scale) and dst.max() < (1 << scale)
    for bit in (scale - 1, scale - 4, 0):
        sb, db = (src >> bit) & 1, (dst >> bit) & 1
        share = [np.mean((sb == i) & (db == j)) for i in (0, 1) for j in (0, 1)]
        assert np.allclose(share, [0.57, 0.19, 0.19, 0.05], atol=0.01)
    # slices of the stream are reproducible independently
    s2, d2 = go.rmat_edgelist(scale, 1000, seed=3, first_edge=5000)
    assert np.array_equal(s2, src[5000:6000]) and np.array_equal(d2, dst[5000:6000])
    # clip_and_flip keeps edges in the lower triangle (generate_rmat_test.cpp)
    s3, d3 = go.rmat_edgelist(scale, 1 << 14, seed=1, clip_and_flip=True)
    assert np.all(s3 >= d3)


@pytest.mark.parametrize("case", ["karate", "rmat", "multi"])
def test_compressed_build_degrees_and_weight_sums_match_the_compiled_reference(oracle_built, case):
    """oracle/_ref holds the reference's own host references for the graph structure
    (graph_test.cpp graph_reference, degree_test.cpp degree_reference, weight_sum_test.cpp
    weight_sum_reference).  The oracle's renumber + compress must describe the same graph: per
    row the same (neighbour, weight) pairs once the labels are mapped back - compared sorted, as
    graph_test.cpp:154-190 does - and the same degrees / weight sums."""
    if not oracle_built.ref_available():
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(12)
    if case == "karate":
        src, dst, w, nv, _ = util.load_mtx("karate.mtx", np.float64)
    elif case == "rmat":
        src, dst = util.rmat(10, seed=4)
        nv = 1 << 10
        w = rng.random(len(src))
    else:  # parallel edges and self loops
        nv = 50
        src = rng.integers(0, nv, 4000).astype(np.int32)
        dst = rng.integers(0, nv, 4000).astype(np.int32)
        w = rng.integers(1, 9, 4000).astype(np.float64)
    verts = np.arange(nv, dtype=np.int32)
    r_off, r_idx, r_w = oracle_built.graph_ref(src, dst, w, nv, store_transposed=True)
    labels, seg, nmaj, nmin = go.renumber_sg(dst.astype(np.int32), src.astype(np.int32), verts)
    offs, idx, ww = go.compress_sorted(nmaj, nmin, w, 0, nv)
    assert offs[-1] == r_off[-1] == len(src)
    for r in range(nv):
        ext = int(labels[r])
        mine = sorted(zip(labels[idx[offs[r]:offs[r + 1]]].tolist(), ww[offs[r]:offs[r + 1]].tolist()))
        ref = sorted(zip(r_idx[r_off[ext]:r_off[ext + 1]].tolist(),
                         r_w[r_off[ext]:r_off[ext + 1]].tolist()))
        assert mine == ref, f"row {ext}"
    # degrees: rows are destinations, so major = in-degree, minor = out-degree
    in_deg = oracle_built.degree_ref(r_off, r_idx, nv, True)
    out_deg = oracle_built.degree_ref(r_off, r_idx, nv, False)
    assert np.array_equal(np.diff(offs), in_deg[labels])
    assert np.array_equal(go.out_weight_sums(idx, None, nv, np.int64), out_deg[labels])
    # out-weight sums (the PageRank denominator): graph_view_t::compute_out_weight_sums
    out_w = oracle_built.weight_sum_ref(r_off, r_idx, r_w, nv, False)
    assert np.allclose(go.out_weight_sums(idx, ww, nv, np.float64), out_w[labels], rtol=1e-12)
    in_w = oracle_built.weight_sum_ref(r_off, r_idx, r_w, nv, True)
    mine_in = np.add.reduceat(np.append(ww, 0.0), offs[:-1]) * (np.diff(offs) > 0)
    assert np.allclose(mine_in, in_w[labels], rtol=1e-12)
