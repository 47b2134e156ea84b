"""Headline benchmark: PageRank GTEPS on a synthetic R-MAT graph (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--scale S] [--dtype f32|f64]
  python bench.py --impl reference ...      # the reference's own host PageRank on CPU cores

A "step" is one PageRank power iteration = one pass of the pull kernel over the whole graph.
One JSON line is printed by rank 0 (contract in the task description; fields documented in
DESIGN.md "measurement").
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

A, B, C = 0.57, 0.19, 0.19          # benchmarks/python_e2e/cugraph_dask_funcs.py:54-65
ALPHA, EPSILON = 0.85, 1e-6         # cpp/tests/link_analysis/pagerank_test.cpp:228-229
EDGE_FACTOR = 16


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scale", type=int, default=24)
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--cpu-scale", type=int, default=None,
                    help="scale of the CPU sample (default: the same graph as the GPU arm)")
    ap.add_argument("--cpu-iterations", type=int, default=3,
                    help="pinned iterations of the cpu_baseline sample of the b200 arm")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


def algorithmic_bytes(nv, ne, val_size, weighted=False, off_size=4):
    """offsets + indices + weights + gathered values per iteration."""
    return (nv + 1) * off_size + ne * 4 + (ne * val_size if weighted else 0) + ne * val_size


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.samples, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "25"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        # median of the upper half = clocks while the kernels were actually running
        load = sm[len(sm) // 2:] if sm else []
        return {"sm_mhz": (load[len(load) // 2] if load else None), "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------
# CPU arm: the reference's own host pagerank_reference (oracle/_ref) on the SAME graph
# ------------------------------------------------------------------------------------------
def workload_string(scale, dtype):
    nv, ne = 1 << scale, EDGE_FACTOR << scale
    return (f"PageRank {dtype} on RMAT scale {scale} ef {EDGE_FACTOR} ({nv} V, {ne} E), "
            f"a/b/c={A}/{B}/{C}, alpha={ALPHA}; a step = one power iteration (one pull-kernel pass "
            "over the graph)")


def cpu_reference_gteps(scale, dtype, iterations, edges=None):
    """Times oracle/_ref (the reference's pagerank_reference compiled from /root/reference; the
    bit-identical port when _ref was not built) for `iterations` pinned power iterations on the
    R-MAT graph of `scale` - the same edge list the GPU arm runs (the generators are
    bit-identical twins).  One thread: the reference's host implementation is scalar code.  The
    call includes pagerank_reference's own out-weight pass over the edges."""
    import numpy as np
    from oracle import cpu_oracle
    cpu_oracle.build()
    ne, nv = EDGE_FACTOR << scale, 1 << scale
    t0 = time.perf_counter()
    src, dst = edges if edges is not None else cpu_oracle.rmat_edgelist_fast(scale, ne, a=A, b=B, c=C, seed=0)
    offsets, indices = cpu_oracle.csc_from_coo_fast(src, dst, nv, threads=8)
    del src, dst
    prep = time.perf_counter() - t0
    npdt = np.float32 if dtype == "f32" else np.float64
    kind = "reference" if cpu_oracle.ref_available() else "port"
    t0 = time.perf_counter()
    cpu_oracle.pagerank(offsets, indices, None, dtype=npdt, alpha=ALPHA, fixed_iterations=iterations,
                        impl="ref" if kind == "reference" else "port")
    dt = time.perf_counter() - t0
    return {"value": ne * iterations / dt / 1e9, "unit": "GTEPS", "cores": 1, "kind": kind,
            "scale": scale, "iterations": iterations, "seconds": dt,
            "sample": f"RMAT scale {scale} ef {EDGE_FACTOR} ({ne} edges, the GPU arm's graph), "
                      f"{dtype}, {iterations} pinned iterations in {dt:.2f} s on 1 of "
                      f"{os.cpu_count()} host cores (input built in {prep:.1f} s, untimed)"}


def bounded_iterations(scale, want, budget_s=120.0, assumed_gteps=0.2):
    """iterations of the scalar CPU code that fit the time budget (it runs ~0.2-0.35 GTEPS)"""
    ne = EDGE_FACTOR << scale
    return max(1, min(want, int(budget_s * assumed_gteps * 1e9 / ne)))


def run_reference(args, rank, world):
    if rank != 0:
        return
    scale = args.scale if args.cpu_scale is None else args.cpu_scale
    iters = bounded_iterations(scale, max(args.steps, 1))
    best = cpu_reference_gteps(scale, args.dtype, iters)
    line = {
        "impl": "reference", "metric": "pagerank_gteps", "value": best["value"], "unit": "GTEPS",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": best["seconds"] / iters * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": workload_string(scale, args.dtype),
                   "same_config": scale == args.scale,
                   "reference_arm": "the reference's host pagerank_reference "
                                    "(cpp/tests/link_analysis/pagerank_test.cpp:42-130) on CPU, "
                                    f"{iters} pinned iterations; see cpu_baseline.sample"},
        "cpu_baseline": best,
        "e2e": {"value": best["value"], "unit": "GTEPS", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def parity_check(engine, h, rank, world, dtype, tdt):
    """Self-check of the path this process group is about to time (SCALE carried no
    correctness evidence): R-MAT scale 20 through the same N-rank graph build + PageRank, 20
    pinned iterations, gathered to rank 0 and compared per vertex with the CPU oracle (fp64 port,
    pinned bit-identical to the compiled reference; extended-precision sums for f64)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    scale, iters = 20, 20
    nv, ne = 1 << scale, EDGE_FACTOR << scale
    per = ne // world
    first = rank * per
    cnt = per if rank < world - 1 else ne - first
    src, dst = engine.generate_rmat_edgelist(h, scale, cnt, a=A, b=B, c=C, seed=0, first_edge=first)
    vper = nv // world
    verts = torch.arange(rank * vper, nv if rank == world - 1 else (rank + 1) * vper,
                         dtype=torch.int32, device="cuda")
    g = engine.Graph(h, src, dst, vertices=verts, multi_gpu=world > 1)
    pr, st = engine.pagerank(g, alpha=ALPHA, epsilon=EPSILON, dtype=tdt, fixed_iterations=iters)
    ids = g.renumber_map().cpu().numpy()
    vals = pr.cpu().numpy()
    del g, src, dst
    if world > 1:
        parts = [None] * world if rank == 0 else None
        dist.gather_object((ids, vals), parts, dst=0)
    else:
        parts = [(ids, vals)]
    if rank != 0:
        return None
    from oracle import cpu_oracle
    cpu_oracle.build()
    ext = np.full(nv, np.nan, dtype=np.float64)
    for i, v in parts:
        ext[i] = v
    es, ed = cpu_oracle.rmat_edgelist_fast(scale, ne, a=A, b=B, c=C, seed=0)
    offsets, indices = cpu_oracle.csc_from_coo_fast(es, ed, nv, threads=8)
    ref, _ = cpu_oracle.pagerank(offsets, indices, None, dtype=np.float64, alpha=ALPHA,
                                 fixed_iterations=iters, impl="ext" if dtype == "f64" else "port")
    err = np.abs(ext - ref) / np.maximum(np.abs(ref), np.finfo(np.float64).tiny)
    tol = 1e-5 if dtype == "f32" else 1e-12
    top = lambda x: np.lexsort((np.arange(nv), -x))[:100]
    complete = bool(not np.isnan(ext).any())
    max_err = float(np.nanmax(err)) if complete else float("inf")
    total = float(np.nansum(ext))
    same_top = bool(np.array_equal(top(ext), top(ref)))
    ok = complete and max_err <= tol and abs(total - 1.0) < (1e-4 if dtype == "f32" else 1e-9) and same_top
    return {"workload": f"PageRank {dtype}, RMAT scale {scale}, {iters} pinned iterations, {world} rank(s)",
            "oracle": "fp64 port of pagerank_reference" + (" with extended-precision sums" if dtype == "f64" else ""),
            "max_rel_err": max_err, "tolerance": tol, "sum": total, "top100_identical": same_top,
            "pass": bool(ok)}


def run_b200(args, rank, world):
    import torch
    import torch.distributed as dist
    from cugraph_forked_b200 import engine

    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local_rank)
    tdt = torch.float32 if args.dtype == "f32" else torch.float64
    val_size = 4 if args.dtype == "f32" else 8
    scale = args.scale
    nv, ne = 1 << scale, EDGE_FACTOR << scale

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        uid = [engine.Handle.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
    if world > 1:
        from cugraph_forked_b200.partition import grid_2d
        rows, cols = grid_2d(world)
        h = engine.Handle.multi_gpu(uid[0], rank, world, rows)
    else:
        h = engine.Handle()

    # ---- self-check of this N-rank path against the CPU oracle, before anything is timed ----
    parity = None
    if not args.no_parity:
        parity = parity_check(engine, h, rank, world, args.dtype, tdt)
        torch.cuda.empty_cache()

    # ---- synthetic input: this rank's slice of the R-MAT edge stream, resident in HBM ----
    per = ne // world
    first = rank * per
    cnt = per if rank < world - 1 else ne - first
    src, dst = engine.generate_rmat_edgelist(h, scale, cnt, a=A, b=B, c=C, seed=0, first_edge=first)
    if world > 1:
        vper = nv // world
        verts = torch.arange(rank * vper, nv if rank == world - 1 else (rank + 1) * vper,
                             dtype=torch.int32, device="cuda")
        g = engine.Graph(h, src, dst, vertices=verts, multi_gpu=True)
    else:
        verts = torch.arange(nv, dtype=torch.int32, device="cuda")
        g = engine.Graph(h, src, dst, vertices=verts)

    # ---- device-resident timing: W warm-up iterations, then exactly K timed iterations ----
    engine.pagerank(g, alpha=ALPHA, epsilon=EPSILON, dtype=tdt, fixed_iterations=max(args.warmup, 3))
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    _, st = engine.pagerank(g, alpha=ALPHA, epsilon=EPSILON, dtype=tdt, fixed_iterations=args.steps)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = torch.tensor([st.elapsed_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    gteps = ne * args.steps / ms_total / 1e6

    # ---- roofline of the dominant kernel (pull_tiles_kernel) from its per-launch events ----
    peak, peak_kind = measured_peak()
    # per-GPU algorithmic bytes: C local partitions of V/P rows, E/P edges
    if world > 1:
        b_iter = cols * (nv // world + 1) * 4 + (ne // world) * (4 + val_size)
    else:
        b_iter = algorithmic_bytes(nv, ne, val_size)
    # one SpMV = phase 1 (pre-reduction of the cells of x's bands from shared memory) + phase 2
    # (rows of the reduced CSC), timed together by the library; + the small tail kernel
    if os.environ.get("B200_BANDED", "1") == "0":
        path, kernels = "tiled", "pull_tiles_kernel"
    elif os.environ.get("B200_LAYOUT", "1" if (world == 1 and ne < (1 << 25)) else "2") != "2":
        # the library's own rule: round 1's banded path on single-GPU graphs below 2^25 edges
        path, kernels = "banded-r1", "band_reduce_kernel + pull_tiles_kernel (one SpMV)"
    else:
        path, kernels = "sell-r2", "sell_band_kernel + sell_rows_kernel (one SpMV)"
    kern_ms = st.spmv_ms / max(st.spmv_launches, 1)
    launches_per_iter = max(st.spmv_launches, 1) / max(min(args.steps, 256), 1)
    achieved = (b_iter / launches_per_iter) / kern_ms / 1e6 if kern_ms > 0 else None
    # dram bytes per SpMV from the committed ncu capture of THIS kernel build: the entry is used
    # only if its build tag equals the tag the loaded library reports, else null
    traffic = None
    build_tag = engine.build_info()
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            ent = json.load(open(tpath)).get(f"s{scale}_{args.dtype}_n{world}_{path}")
            if ent and ent.get("build") == build_tag:
                traffic = ent.get("bytes")
        except Exception:
            traffic = None

    # ---- end to end through the public API with HOST buffers ----
    e2e = None
    if not args.no_e2e:
        hs, hd = src.cpu().pin_memory(), dst.cpu().pin_memory()
        hv = verts.cpu().pin_memory()
        n_local = g.num_local_vertices
        # pinned destination buffers, like the pinned sources, exist before the timed region
        ids = torch.empty(n_local, dtype=torch.int32, pin_memory=True)
        out = torch.empty(n_local, dtype=tdt, pin_memory=True)
        del g, src, dst
        torch.cuda.empty_cache()
        h2d = (hs.numel() + hd.numel() + hv.numel()) * 4
        d2h = out.numel() * val_size + ids.numel() * 4
        # the whole end-to-end call is made twice and the faster one reported: a single call was
        # seen to take 7x longer once on a fresh box (a host-side hiccup in the 2.2 GB upload)
        runs = []
        for attempt in range(2):
            barrier()
            t0 = time.perf_counter()
            ds, dd, dv = hs.cuda(non_blocking=True), hd.cuda(non_blocking=True), hv.cuda(non_blocking=True)
            torch.cuda.synchronize()          # the graph build would wait for the copies anyway
            t_h2d = time.perf_counter()
            g2 = engine.Graph(h, ds, dd, vertices=dv, multi_gpu=world > 1)
            t_build = time.perf_counter()
            pr, st2 = engine.pagerank(g2, alpha=ALPHA, epsilon=EPSILON, dtype=tdt,
                                      fixed_iterations=args.steps)
            t_run = time.perf_counter()
            # results land in pinned host memory (as the inputs come from pinned memory)
            ids_d = g2.renumber_map()
            ids.copy_(ids_d, non_blocking=True)
            out.copy_(pr, non_blocking=True)
            torch.cuda.synchronize()
            t_end = time.perf_counter()
            dt = torch.tensor([t_end - t0], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            runs.append((float(dt.item()), {"h2d": (t_h2d - t0) * 1e3, "graph_build": (t_build - t_h2d) * 1e3,
                                            "layout_build_and_iterations": (t_run - t_build) * 1e3,
                                            "d2h": (t_end - t_run) * 1e3}))
            del g2, ds, dd, dv, pr, ids_d
            torch.cuda.empty_cache()
        best = min(runs, key=lambda r: r[0])
        e2e = {"value": ne * args.steps / best[0] / 1e9, "unit": "GTEPS",
               "h2d_bytes_per_step": h2d // args.steps, "d2h_bytes_per_step": d2h // args.steps,
               "what": "host COO (pinned) -> H2D -> graph build -> layout build -> K iterations -> "
                       "D2H of (vertex ids, scores), wall clock, max over ranks; best of 2 complete "
                       "calls; bytes are per-call totals / K",
               "seconds_of_both_calls": [r[0] for r in runs], "rank0_ms": best[1]}

    # ---- collectives of one iteration: allgatherv over row_comm lands
    #      (R-1)/R * V/C values per GPU, the per-partition reductions over col_comm send
    #      (C-1)/C * V/R values per GPU; fraction of one NVLink direction (900 GB/s) they would
    #      need if they were not overlapped with the kernels ----
    nvlink = None
    if world > 1:
        recv_b = (rows - 1) / rows * (nv / cols) * val_size
        send_b = (cols - 1) / cols * (nv / rows) * val_size
        step_s = ms_total / args.steps / 1e3
        nvlink = {"allgather_recv_bytes_per_gpu": int(recv_b), "reduce_send_bytes_per_gpu": int(send_b),
                  "frac_of_900GBps_per_direction": max(recv_b, send_b) / step_s / 900e9}
    if rank != 0:
        return
    cpu = None
    if not args.no_cpu:
        cs = scale if args.cpu_scale is None else args.cpu_scale
        cpu = cpu_reference_gteps(cs, args.dtype, bounded_iterations(cs, args.cpu_iterations, 40.0))
        cpu["same_config"] = cs == scale
    line = {
        "metric": "pagerank_gteps", "value": gteps, "unit": "GTEPS", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": args.dtype,
        "data": "synthetic",
        "config": {"workload": workload_string(scale, args.dtype),
                   "l2": "inputs larger than L2 (no flush between iterations)" if b_iter > 126e6
                         else "working set fits L2; no flush (dram bytes reported in profiles/)",
                   "partition": "single GPU" if world == 1 else f"2D {rows}x{cols} (row x col comm)",
                   "pull_path": path, "library": build_tag},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": (achieved / peak) if achieved else None, "traffic": traffic,
                     "peak_source": peak_kind,
                     "kernel": kernels,
                     "algorithmic_bytes_per_launch": b_iter / launches_per_iter,
                     "kernel_ms_avg": kern_ms, "kernel_share_of_step":
                         (st.spmv_ms / (ms_total * min(args.steps, 256) / args.steps))
                         if ms_total > 0 else None},
        "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": st.gpu_launches, "clocks": clocks,
        "parity_check": parity,
    }
    if nvlink:
        line["nvlink"] = nvlink
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_b200(args, rank, world)
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
