"""CPU, world sizes 2 and 4 over gloo: the multi-GPU exchange pattern on the HOST side.

Each process plays one rank: its share of the graph comes from the virtual-rank oracle
(`build_mg`), its ranges and communicator memberships from the product's
`cugraph_forked_b200.partition`, the per-partition SpMV is numpy; the collectives are the ones
csrc/mg.cu issues over NCCL (all-gather of the local slice over the row communicator, reduce of
every partition's partial sums onto its owner over the col communicator, 2-scalar all-reduce).
The result must equal the single-process oracle PageRank - which is what
cpp/tests/link_analysis/mg_pagerank_test.cpp:175-270 asserts for the reference (MG == SG)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, scale, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from cugraph_forked_b200 import partition
    from oracle import graph_oracle as go

    src, dst = go.rmat_edgelist(scale, 8 << scale, seed=3)
    rng = np.random.default_rng(1)
    wgt = rng.random(len(src)) + 0.1
    mg = go.build_mg(src, dst, wgt, world)
    part = partition.Partition(mg["vertex_partition_offsets"], world, rank)
    R, C = part.row_comm_size, part.col_comm_size
    assert (R, C) == (mg["R"], mg["C"])
    row_ranks = partition.row_comm_ranks(rank, world)
    col_ranks = partition.col_comm_ranks(rank, world)
    assert row_ranks[part.row_comm_rank] == rank and col_ranks[part.col_comm_rank] == rank
    # every rank creates every group, in the same order (torch.distributed rule)
    row_groups = {tuple(partition.row_comm_ranks(r, world)) for r in range(world)}
    col_groups = {tuple(partition.col_comm_ranks(r, world)) for r in range(world)}
    groups = {g: dist.new_group(list(g)) for g in sorted(row_groups | col_groups)}
    row_group, col_group = groups[tuple(row_ranks)], groups[tuple(col_ranks)]

    mine = mg["parts"][rank]
    lo, hi = part.local_vertex_partition_range()
    nloc, nv = hi - lo, mg["num_vertices"]
    minor_first, minor_last = part.local_edge_partition_minor_range()
    for i in range(C):
        assert (mine[i]["major_first"], mine[i]["major_last"]) == part.local_edge_partition_major_range(i)
        assert (mine[i]["minor_first"], mine[i]["minor_last"]) == (minor_first, minor_last)

    # out-weight sums of the local vertices: histogram over the minor range, reduced over the row
    hist = np.zeros(minor_last - minor_first)
    for p in mine:
        np.add.at(hist, p["indices"] - minor_first, p["weights"])
    t = torch.from_numpy(hist)
    dist.all_reduce(t, group=row_group)
    outw = t.numpy()[lo - minor_first:hi - minor_first].copy()

    alpha, eps = 0.85, 1e-10
    pr = np.full(nloc, 1.0 / nv)
    sizes = [mg["vertex_partition_offsets"][r + 1] - mg["vertex_partition_offsets"][r] for r in row_ranks]
    iters = 0
    for it in range(500):
        x_local = np.where(outw > 0, pr / np.where(outw > 0, outw, 1.0), 0.0)
        scal = torch.tensor([pr[outw == 0].sum(), 0.0], dtype=torch.float64)
        dist.all_reduce(scal)
        dangling = float(scal[0])
        # 1. allgatherv over the row communicator -> the minor-range buffer
        pieces = [torch.zeros(int(s), dtype=torch.float64) for s in sizes]
        if R > 1:
            # gloo all_gather needs equal sizes: pad to the largest piece
            m = max(int(s) for s in sizes)
            pad = torch.zeros(m, dtype=torch.float64)
            pad[:nloc] = torch.from_numpy(x_local)
            got = [torch.zeros(m, dtype=torch.float64) for _ in sizes]
            dist.all_gather(got, pad, group=row_group)
            pieces = [g[:int(s)] for g, s in zip(got, sizes)]
        else:
            pieces = [torch.from_numpy(x_local)]
        x_minor = torch.cat(pieces).numpy()
        # 2. every local partition, then reduce onto the col_comm rank that owns its majors
        new_sum = None
        for i in range(C):
            p = mine[i]
            contrib = x_minor[p["indices"] - minor_first] * p["weights"]
            part_sum = np.add.reduceat(np.concatenate([contrib, [0.0]]),
                                       np.minimum(p["offsets"][:-1], len(contrib)))
            part_sum[np.diff(p["offsets"]) == 0] = 0.0
            tt = torch.from_numpy(np.ascontiguousarray(part_sum))
            dist.reduce(tt, dst=col_ranks[i], group=col_group)
            if col_ranks[i] == rank:
                new_sum = tt.numpy()
        # 3. local epilogue + the two scalars
        new_pr = alpha * new_sum + (dangling * alpha + (1 - alpha)) / nv
        scal = torch.tensor([0.0, np.abs(new_pr - pr).sum()], dtype=torch.float64)
        dist.all_reduce(scal)
        pr, iters = new_pr, it + 1
        if float(scal[1]) < eps:
            break
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), labels=mg["labels"][rank], pr=pr, iters=iters)
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_mg_exchange_pattern_over_gloo(world, tmp_path, oracle_built):
    from oracle import graph_oracle as go
    from tests import util
    scale = 9
    port = _free_port()
    mp.spawn(_rank_main, args=(world, port, scale, str(tmp_path)), nprocs=world, join=True)
    src, dst = go.rmat_edgelist(scale, 8 << scale, seed=3)
    wgt = np.random.default_rng(1).random(len(src)) + 0.1
    ext = np.unique(np.concatenate([src, dst]))
    relabel = np.full(int(ext.max()) + 1, -1, np.int64)
    relabel[ext] = np.arange(len(ext))
    ref, ref_iters = util.oracle_pagerank(relabel[src].astype(np.int32), relabel[dst].astype(np.int32),
                                          wgt, len(ext), dtype=np.float64, alpha=0.85, epsilon=1e-10,
                                          max_iterations=500)
    seen = 0
    for r in range(world):
        z = np.load(tmp_path / f"rank{r}.npz")
        assert int(z["iters"]) == ref_iters
        got = z["pr"]
        want = ref[relabel[z["labels"]]]
        assert util.rel_err(got, want).max() <= 1e-10
        seen += len(got)
    assert seen == len(ext)


def test_partition_ranges_match_the_oracle():
    from cugraph_forked_b200 import partition
    from oracle import graph_oracle as go
    for world in (1, 2, 4, 8):
        assert partition.grid_2d(world) == go.grid_2d(world)
        R, C = partition.grid_2d(world)
        for rank in range(world):
            assert partition.rank_coords(rank, R) == go.rank_coords(rank, R)
            rr, cr = partition.rank_coords(rank, R)
            assert partition.row_comm_ranks(rank, world) == [cr * R + j for j in range(R)]
            assert partition.col_comm_ranks(rank, world) == [i * R + rr for i in range(C)]
    src, dst = go.rmat_edgelist(8, 8 << 8, seed=5)
    mg = go.build_mg(src, dst, None, 8)
    for rank in range(8):
        p = partition.Partition(mg["vertex_partition_offsets"], 8, rank)
        for i in range(p.number_of_local_edge_partitions()):
            want = mg["parts"][rank][i]
            assert p.local_edge_partition_major_range(i) == (want["major_first"], want["major_last"])
            assert p.local_edge_partition_minor_range() == (want["minor_first"], want["minor_last"])
