"""Multi-GPU parity worker: one process per GPU (torchrun) or in-process with world size 1.

Every rank builds the SAME synthetic edge list on the host, feeds its contiguous slice to
b200_graph_create_mg and compares what it holds afterwards with the virtual-rank oracle
(oracle/graph_oracle.py build_mg): vertex partition ranges, renumber labels, segment offsets and
every local edge partition's offsets / indices bit-exactly, weights exactly (stable order); then
PageRank / Katz / eigenvector scores against the CPU oracle run on the whole graph, as
cpp/tests/link_analysis/mg_pagerank_test.cpp:175-270 compares MG with SG results.

Run:  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 \
          --master-port 29511 tests/mg_worker.py [--scale 12]
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def make_handle():
    """(handle, rank, world): NCCL unique id from rank 0 travels over a gloo broadcast."""
    from cugraph_forked_b200 import engine, partition
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        if not dist.is_initialized():
            dist.init_process_group("gloo")
        box = [engine.Handle.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        uid = box[0]
    else:
        uid = engine.Handle.nccl_unique_id()
    r, _ = partition.grid_2d(world)
    return engine.Handle.multi_gpu(uid, rank, world, r), rank, world


def check_graph(g, oracle, rank, weighted):
    """bit-exact comparison of this rank's share of the graph with the oracle's"""
    from oracle import graph_oracle as go
    R, C = oracle["R"], oracle["C"]
    assert g.vertex_partition_offsets() == [int(x) for x in oracle["vertex_partition_offsets"]]
    assert g.num_vertices == oracle["num_vertices"]
    labels = g.renumber_map().cpu().numpy()
    assert np.array_equal(labels, oracle["labels"][rank]), "renumber labels differ"
    assert g.segment_offsets() == [int(x) for x in oracle["segment_offsets"][rank]]
    row_rank, col_rank = go.rank_coords(rank, R)
    for i in range(C):
        want = oracle["parts"][rank][i]
        v = g.view(i)
        assert (v.major_first, v.major_last) == (want["major_first"], want["major_last"])
        assert (v.minor_first, v.minor_last) == (want["minor_first"], want["minor_last"])
        assert g.partition_segment_offsets(i) == [int(x) for x in want["segment_offsets"]]
        off = g.offsets(i).cpu().numpy().astype(np.int64)
        assert np.array_equal(off, want["offsets"]), f"offsets of partition {i} differ"
        idx = g.indices(i).cpu().numpy().astype(np.int64) + v.minor_first
        assert np.array_equal(idx, want["indices"]), f"indices of partition {i} differ"
        if weighted:
            w = g.weights(i).cpu().numpy()
            assert np.array_equal(w, want["weights"]), f"weights of partition {i} differ"
        # DCSC form of the hypersparse segment (compress_hypersparse_offsets,
        # structure_utils.cuh:109-159): present exactly when the oracle says so, bit-exact
        dcs_want = go.compress_hypersparse(want["offsets"], want["major_first"],
                                           [int(x) for x in want["segment_offsets"]])
        dcs = g.dcs(i)
        assert (dcs is None) == (dcs_want is None), f"DCSC presence of partition {i} differs"
        if dcs is not None:
            assert np.array_equal(dcs[0].cpu().numpy().astype(np.int64), dcs_want[0]), \
                f"dcs offsets of partition {i} differ"
            assert np.array_equal(dcs[1].cpu().numpy().astype(np.int64), dcs_want[1]), \
                f"dcs_nzd_vertices of partition {i} differ"
            assert dcs[2] == want["major_first"] + int(want["segment_offsets"][3])


def run(scale: int, edge_factor: int, dtype, weighted: bool, handle, rank, world, verbose=True):
    from cugraph_forked_b200 import engine
    from oracle import cpu_oracle
    from oracle import graph_oracle as go
    from tests import util

    ne = edge_factor << scale
    src, dst = go.rmat_edgelist(scale, ne, seed=7)
    rng = np.random.default_rng(11)
    wgt = (rng.random(ne) + 0.05).astype(dtype) if weighted else None
    lo, hi = rank * ne // world, (rank + 1) * ne // world
    t = lambda a: None if a is None else torch.as_tensor(a[lo:hi]).cuda()
    g = engine.Graph(handle, t(src), t(dst), t(wgt), multi_gpu=True)
    oracle = go.build_mg(src, dst, wgt, world)
    check_graph(g, oracle, rank, weighted)

    labels = oracle["labels"][rank]
    ext_all = np.unique(np.concatenate([src, dst]))
    relabel = np.full(int(ext_all.max()) + 1, -1, np.int64)
    relabel[ext_all] = np.arange(len(ext_all))
    nv = len(ext_all)
    s2, d2 = relabel[src].astype(np.int32), relabel[dst].astype(np.int32)
    mine = relabel[labels]
    tdtype = torch.float32 if dtype == np.float32 else torch.float64
    tol = 1e-5 if dtype == np.float32 else 1e-11

    # out-weight sums
    ow = g.out_weight_sums(torch.float64).cpu().numpy()
    ow_ref = np.zeros(nv)
    np.add.at(ow_ref, s2, 1.0 if wgt is None else wgt.astype(np.float64))
    assert np.allclose(ow, ow_ref[mine], rtol=1e-6 if weighted else 0, atol=0)

    # PageRank at a pinned iteration count against the fp64 oracle
    pr, st = engine.pagerank(g, alpha=0.85, epsilon=1e-6 if dtype == np.float32 else 1e-12,
                             max_iterations=500, dtype=tdtype)
    assert st.converged
    ref, _ = util.oracle_pagerank(s2, d2, wgt, nv, dtype=np.float64,
                                  fixed_iterations=st.iterations)
    err = util.rel_err(pr.cpu().numpy(), ref[mine]).max() if len(mine) else 0.0
    assert err <= tol, f"MG PageRank differs: {err:.3e}"

    # Katz
    kz, stk = engine.katz_centrality(g, alpha=0.001, beta=1.0, epsilon=1e-7, max_iterations=500,
                                     dtype=tdtype, normalize=True)
    refk, _ = util.oracle_katz(s2, d2, wgt, nv, dtype=np.float64, alpha=0.001, beta=1.0,
                               normalize=True, fixed_iterations=stk.iterations)
    errk = util.rel_err(kz.cpu().numpy(), refk[mine]).max() if len(mine) else 0.0
    assert errk <= tol, f"MG Katz differs: {errk:.3e}"

    # eigenvector (fixed iteration count: the power iteration oscillates on a directed R-MAT)
    ev, ste = engine.eigenvector_centrality(g, epsilon=1e-6, max_iterations=500, dtype=tdtype,
                                            fixed_iterations=20)
    refe, _ = cpu_oracle.eigenvector(d2, s2, wgt, num_vertices=nv, dtype=np.float64,
                                     fixed_iterations=20)
    erre = (np.abs(ev.cpu().numpy() - refe[mine]).max() / refe.max()) if len(mine) else 0.0
    assert erre <= (1e-4 if dtype == np.float32 else 1e-10), f"MG eigenvector differs: {erre:.3e}"

    # personalized PageRank: every rank personalises a few of its own vertices
    if len(mine) >= 4:
        first = g.view().local_vertex_first
        pv_local = np.arange(0, min(len(mine), 16), 2, dtype=np.int32)
        pw = (np.arange(len(pv_local)) + 1.0).astype(dtype)
    else:
        first, pv_local, pw = 0, np.zeros(0, np.int32), np.zeros(0, dtype)
    ppr, stp = engine.pagerank(g, alpha=0.85, epsilon=1e-6 if dtype == np.float32 else 1e-12,
                               max_iterations=500, dtype=tdtype,
                               personalization_vertices=torch.as_tensor(pv_local + first).cuda(),
                               personalization_values=torch.as_tensor(pw).cuda())
    # the oracle needs the union of every rank's pairs
    pv_all, pw_all = [], []
    for r in range(world):
        m = relabel[oracle["labels"][r]]
        if len(m) >= 4:
            sel = np.arange(0, min(len(m), 16), 2)
            pv_all.append(m[sel])
            pw_all.append(np.arange(len(sel)) + 1.0)
    pv_all = np.concatenate(pv_all).astype(np.int32)
    pw_all = np.concatenate(pw_all)
    refp, _ = util.oracle_pagerank(s2, d2, wgt, nv, dtype=np.float64,
                                   personalization=(pv_all, pw_all),
                                   fixed_iterations=stp.iterations)
    errp = util.rel_err(ppr.cpu().numpy(), refp[mine]).max() if len(mine) else 0.0
    assert errp <= tol * 5, f"MG personalized PageRank differs: {errp:.3e}"
    if verbose:
        print(f"MG_WORKER_OK rank={rank}/{world} scale={scale} dtype={np.dtype(dtype).name} "
              f"weighted={weighted} V={nv} iters={st.iterations} pr_err={err:.2e} "
              f"katz_err={errk:.2e} eig_err={erre:.2e} ppr_err={errp:.2e}", flush=True)


def run_tiny(handle, rank, world):
    """The 6-vertex golden graph of c_api/mg_pagerank_test.c:213-247 on any number of ranks: most
    ranks own one vertex or none and hold no edges at all (empty partitions, empty exchanges)."""
    from cugraph_forked_b200 import engine
    from oracle import graph_oracle as go
    from tests import golden_vectors as gv
    src, dst, wgt = gv.G6_SRC, gv.G6_DST, gv.G6_WGT
    # every edge enters through rank 0; the other ranks contribute nothing
    sel = slice(0, len(src)) if rank == 0 else slice(0, 0)
    t = lambda a: torch.as_tensor(a[sel]).cuda()
    g = engine.Graph(handle, t(src), t(dst), t(wgt), multi_gpu=True)
    oracle = go.build_mg(src, dst, wgt, world)
    check_graph(g, oracle, rank, True)
    p = gv.G6_PAGERANK
    pr, st = engine.pagerank(g, alpha=p["alpha"], epsilon=p["epsilon"],
                             max_iterations=p["max_iterations"])
    labels = oracle["labels"][rank]
    assert gv.nearly_equal(pr.cpu().numpy(), p["result"][labels])
    k = gv.G6_KATZ
    kz, _ = engine.katz_centrality(g, alpha=k["alpha"], beta=k["beta"], epsilon=k["epsilon"],
                                   max_iterations=k["max_iterations"], normalize=True)
    assert gv.nearly_equal(kz.cpu().numpy(), k["result"][labels])
    print(f"MG_WORKER_OK rank={rank}/{world} tiny graph: local vertices {len(labels)}", flush=True)


def run_c_api(handle, rank, world, scale=10):
    """cugraph_personalized_pagerank / nstart / precomputed sums through the C API on a
    multi-GPU graph, with (vertex, value) pairs that belong to OTHER ranks: the library has to
    move them to their owners (c_api.cu handed local offsets to the MG driver and
    rejected foreign vertices).  Every rank passes a disjoint slice of the pairs, chosen by
    position and not by ownership."""
    from cugraph_forked_b200 import pylibcugraph as plc
    from tests import util
    nv, ne = 1 << scale, 16 << scale
    src, dst = util.rmat(scale, seed=21)
    # vertex ids appear in the graph only if they have an edge: restrict the pairs to those
    present = np.zeros(nv, bool)
    present[src] = True
    present[dst] = True
    ids = np.flatnonzero(present).astype(np.int32)
    relabel = np.cumsum(present) - 1
    per = ne // world
    lo, hi = rank * per, (ne if rank == world - 1 else (rank + 1) * per)
    t = lambda a: torch.as_tensor(np.ascontiguousarray(a)).cuda()
    rh = plc.ResourceHandle(handle)
    g = plc.MGGraph(rh, plc.GraphProperties(is_multigraph=True), t(src[lo:hi]), t(dst[lo:hi]),
                    t(np.ones(hi - lo, np.float64)), store_transposed=True, num_edges=hi - lo)
    rng = np.random.default_rng(4)
    pv = rng.permutation(ids)[: len(ids) // 3]
    pw = rng.random(len(pv)) + 0.1
    outw = np.bincount(src, minlength=nv).astype(np.float64)
    guess = rng.random(nv) + 0.5
    mine = slice(rank, None, world)          # this rank's share of every pair list
    verts, vals = plc.personalized_pagerank(
        rh, g, t(ids[mine]), t(outw[ids][mine]), t(ids[mine]), t(guess[ids][mine]), t(pv[mine]),
        t(pw[mine]), 0.85, 1e-12, 500, False)
    s2, d2 = relabel[src].astype(np.int32), relabel[dst].astype(np.int32)
    ref, _ = util.oracle_pagerank(s2, d2, None, len(ids), dtype=np.float64, epsilon=1e-12,
                                  personalization=(relabel[pv].astype(np.int32), pw),
                                  initial_guess=guess[ids], max_iterations=500)
    got = vals.cpu().numpy()
    err = util.rel_err(got, ref[relabel[verts.cpu().numpy()]]).max() if len(got) else 0.0
    assert err <= 1e-8, f"MG C-API personalized PageRank differs: {err:.3e}"
    # an id no rank owns is rejected on every rank
    bad = np.array([nv + 7], np.int32) if rank == 0 else np.zeros(0, np.int32)
    try:
        plc.personalized_pagerank(rh, g, None, None, None, None, t(bad), t(np.ones(len(bad))), 0.85,
                                  1e-6, 100, False)
        raise AssertionError("a vertex outside the graph was accepted")
    except (RuntimeError, ValueError) as e:
        assert "not in the graph" in str(e)
    print(f"MG_WORKER_OK rank={rank}/{world} C API pairs shuffled to owners: err={err:.2e}", flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=int, default=12)
    ap.add_argument("--edge-factor", type=int, default=16)
    args = ap.parse_args()
    from oracle import cpu_oracle
    handle, rank, world = make_handle()
    if rank == 0:
        cpu_oracle.build()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    for dtype, weighted in ((np.float32, False), (np.float64, True), (np.float32, True)):
        run(args.scale, args.edge_factor, dtype, weighted, handle, rank, world)
    run_tiny(handle, rank, world)
    run_c_api(handle, rank, world)
    # a graph with long rows in every partition (hubs) and many empty rows
    run(args.scale + 2, 8, np.float32, False, handle, rank, world)


if __name__ == "__main__":
    main()
