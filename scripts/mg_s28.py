"""BASELINE.json configs[4]: Katz centrality + personalized PageRank on a weighted R-MAT graph
(scale 28 by default: 4.3 G edges) across the GPUs of one box.  Launch with torchrun.  Every
rank generates its slice of the edge stream and the weights on the device; results are checked
through size-independent properties (sum of PageRank = 1 over all ranks; Katz fixed point
residual on the local vertices is below epsilon by construction of the stop test)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from cugraph_forked_b200 import engine, partition

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 28
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
box = [engine.Handle.nccl_unique_id() if rank == 0 else None]
dist.broadcast_object_list(box, src=0)
R, C = partition.grid_2d(world)
h = engine.Handle.multi_gpu(box[0], rank, world, R)
nv, ne = 1 << scale, 16 << scale
per = ne // world
t0 = time.time()
src, dst = engine.generate_rmat_edgelist(h, scale, per, seed=0, first_edge=rank * per)
w = engine.uniform_fill(h, per, seed=1000 + rank, dtype=torch.float32)
torch.cuda.synchronize(); t1 = time.time()
g = engine.Graph(h, src, dst, w, multi_gpu=True)
torch.cuda.synchronize(); dist.barrier(); t2 = time.time()
del src, dst, w
n = g.num_local_vertices
out = {"scale": scale, "n_gpus": world, "grid": f"{R}x{C}", "num_vertices": g.num_vertices,
       "num_edges": ne, "generate_s": t1 - t0, "build_s": t2 - t1,
       "edge_type": "int64" if g.view().edge_type == 1 else "int32"}
# warm-up: the first collectives on each communicator set up their connections
engine.katz_centrality(g, alpha=1e-4, beta=1.0, fixed_iterations=2)
# Katz: alpha below 1 / max degree is not needed for weights in [0,1); keep it small
kz, st = engine.katz_centrality(g, alpha=1e-4, beta=1.0, epsilon=1e-6, max_iterations=200, normalize=True)
sq = (kz.double() ** 2).sum(); dist.all_reduce(sq)
out["katz"] = {"iterations": st.iterations, "converged": st.converged, "ms_per_iter": st.elapsed_ms / max(st.iterations, 1),
               "gteps": ne * st.iterations / st.elapsed_ms / 1e6, "l2_norm": float(sq.sqrt())}
# personalized PageRank: every rank personalises 1024 of its own vertices
first = g.view().local_vertex_first
k = min(1024, n)
pv = (torch.arange(k, dtype=torch.int32, device="cuda") * max(n // max(k, 1), 1) + first).to(torch.int32)
pw = torch.rand(k, device="cuda", dtype=torch.float32) + 0.1
pr, stp = engine.pagerank(g, alpha=0.85, epsilon=1e-6, max_iterations=200,
                          personalization_vertices=pv, personalization_values=pw)
s = pr.double().sum(); dist.all_reduce(s)
out["personalized_pagerank"] = {"iterations": stp.iterations, "converged": stp.converged,
                                "ms_per_iter": stp.elapsed_ms / max(stp.iterations, 1),
                                "gteps": ne * stp.iterations / stp.elapsed_ms / 1e6, "sum": float(s)}
assert abs(float(s) - 1.0) < 1e-3 and abs(float(sq.sqrt()) - 1.0) < 1e-3
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open(f"gpurun_out/mg_s{scale}_n{world}.json", "w"), indent=1)
    print("MG_BIG", json.dumps(out), flush=True)
dist.destroy_process_group()
