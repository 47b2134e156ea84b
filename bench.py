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
    ap.add_argument("--cpu-scale", type=int, default=20, help="scale of the bounded CPU sample")
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
# CPU arm: the reference's own host pagerank_reference (oracle/_ref) on a bounded sample
# ------------------------------------------------------------------------------------------
def cpu_reference_gteps(scale, dtype):
    """Times oracle/_ref (the reference's pagerank_reference compiled from /root/reference), or
    the bit-identical port when _ref was not built, on RMAT `scale` to convergence.  Single
    thread: the reference's host implementation is scalar code."""
    import numpy as np
    from oracle import cpu_oracle, graph_oracle
    cpu_oracle.build()
    ne = EDGE_FACTOR << scale
    nv = 1 << scale
    src, dst = graph_oracle.rmat_edgelist(scale, ne, a=A, b=B, c=C, seed=0)
    offsets, indices, _ = cpu_oracle.csc_from_coo(src, dst, None, nv)
    npdt = np.float32 if dtype == "f32" else np.float64
    kind = "reference" if cpu_oracle.ref_available() else "port"
    t0 = time.perf_counter()
    _, it = cpu_oracle.pagerank(offsets, indices, None, dtype=npdt, alpha=ALPHA, epsilon=EPSILON,
                                max_iterations=10000, impl="ref" if kind == "reference" else "port")
    dt = time.perf_counter() - t0
    if kind == "reference":
        # the shim cannot see the iteration count; the port is bit-identical (tests/test_oracle.py)
        _, it = cpu_oracle.pagerank(offsets, indices, None, dtype=npdt, alpha=ALPHA,
                                    epsilon=EPSILON, max_iterations=10000, impl="port")
    return {"value": ne * it / dt / 1e9, "unit": "GTEPS", "cores": 1, "kind": kind,
            "sample": f"RMAT scale {scale} ef {EDGE_FACTOR} ({ne} edges), {dtype}, {it} iterations "
                      f"to epsilon {EPSILON:g} in {dt:.2f} s on 1 of {os.cpu_count()} host cores"}


def run_reference(args, rank, world):
    if rank != 0:
        return
    scale = min(args.cpu_scale, args.scale)
    vals = []
    for _ in range(max(1, min(args.steps, 1))):  # one bounded sample per run (tens of seconds)
        vals.append(cpu_reference_gteps(scale, args.dtype))
    best = max(vals, key=lambda v: v["value"])
    nv, ne = 1 << args.scale, EDGE_FACTOR << args.scale
    line = {
        "impl": "reference", "metric": "pagerank_gteps", "value": best["value"], "unit": "GTEPS",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": None, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": f"PageRank {args.dtype} on RMAT scale {args.scale} ef {EDGE_FACTOR} "
                               f"({nv} V, {ne} E), a/b/c={A}/{B}/{C}, alpha={ALPHA}",
                   "reference_arm": "the reference's host pagerank_reference "
                                    "(cpp/tests/link_analysis/pagerank_test.cpp:42-130) on CPU; "
                                    "bounded sample, see cpu_baseline.sample"},
        "cpu_baseline": best,
        "e2e": {"value": best["value"], "unit": "GTEPS", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


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
    # the single-GPU SpMV is two launches when the banded pre-reduction is on (band_reduce_kernel
    # + pull_tiles_kernel, timed together as one SpMV) and one otherwise; + the tail kernel
    banded = os.environ.get("B200_BANDED", "1") != "0" and (
        world > 1 or st.gpu_launches >= 3 * max(st.iterations, 1))
    kern_ms = st.spmv_ms / max(st.spmv_launches, 1)
    launches_per_iter = max(st.spmv_launches, 1) / max(min(args.steps, 256), 1)
    achieved = (b_iter / launches_per_iter) / kern_ms / 1e6 if kern_ms > 0 else None
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(
                f"s{scale}_{args.dtype}_n{world}" + ("_banded" if banded else ""))
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
        barrier()
        t0 = time.perf_counter()
        ds, dd, dv = hs.cuda(non_blocking=True), hd.cuda(non_blocking=True), hv.cuda(non_blocking=True)
        g2 = engine.Graph(h, ds, dd, vertices=dv, multi_gpu=world > 1)
        pr, st2 = engine.pagerank(g2, alpha=ALPHA, epsilon=EPSILON, dtype=tdt,
                                  fixed_iterations=args.steps)
        # results land in pinned host memory (as the inputs come from pinned memory)
        ids_d = g2.renumber_map()
        ids.copy_(ids_d, non_blocking=True)
        out.copy_(pr, non_blocking=True)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        h2d = (hs.numel() + hd.numel() + hv.numel()) * 4
        d2h = out.numel() * val_size + ids.numel() * 4
        e2e = {"value": ne * args.steps / float(dt.item()) / 1e9, "unit": "GTEPS",
               "h2d_bytes_per_step": h2d // args.steps, "d2h_bytes_per_step": d2h // args.steps,
               "what": "host COO (pinned) -> H2D -> graph build -> K iterations -> D2H of "
                       "(vertex ids, scores); bytes are per-call totals / K"}

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
    if world == 1 and not args.no_cpu:
        cpu = cpu_reference_gteps(min(args.cpu_scale, scale), args.dtype)
    line = {
        "metric": "pagerank_gteps", "value": gteps, "unit": "GTEPS", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": args.dtype,
        "data": "synthetic",
        "config": {"workload": f"PageRank {args.dtype} on RMAT scale {scale} ef {EDGE_FACTOR} "
                               f"({nv} V, {ne} E), a/b/c={A}/{B}/{C}, alpha={ALPHA}; "
                               "a step = one power iteration (one pull-kernel pass over the graph)",
                   "l2": "inputs larger than L2 (no flush between iterations)" if b_iter > 126e6
                         else "working set fits L2; no flush (dram bytes reported in profiles/)",
                   "partition": "single GPU" if world == 1 else f"2D {rows}x{cols} (row x col comm)",
                   "pull_path": "banded pre-reduction + tiled pull" if banded else "tiled pull"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": (achieved / peak) if achieved else None, "traffic": traffic,
                     "peak_source": peak_kind,
                     "kernel": "band_reduce_kernel + pull_tiles_kernel (one SpMV)" if banded
                               else "pull_tiles_kernel",
                     "algorithmic_bytes_per_launch": b_iter / launches_per_iter,
                     "kernel_ms_avg": kern_ms, "kernel_share_of_step":
                         (st.spmv_ms / (ms_total * min(args.steps, 256) / args.steps))
                         if ms_total > 0 else None},
        "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": st.gpu_launches, "clocks": clocks,
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
