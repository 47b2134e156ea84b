"""Development aid: where the end-to-end time of bench.py's e2e leg goes."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cugraph_forked_b200 import engine

scale, K = int(sys.argv[1]) if len(sys.argv) > 1 else 24, 100
nv, ne = 1 << scale, 16 << scale
h = engine.Handle()
src, dst = engine.generate_rmat_edgelist(h, scale, ne)
verts = torch.arange(nv, dtype=torch.int32, device="cuda")
hs, hd, hv = src.cpu().pin_memory(), dst.cpu().pin_memory(), verts.cpu().pin_memory()
del src, dst, verts
torch.cuda.empty_cache()
for rep in range(3):
    torch.cuda.synchronize()
    t = [time.perf_counter()]
    def mark():
        torch.cuda.synchronize(); t.append(time.perf_counter())
    ds, dd, dv = hs.cuda(non_blocking=True), hd.cuda(non_blocking=True), hv.cuda(non_blocking=True)
    mark()
    g = engine.Graph(h, ds, dd, vertices=dv)
    mark()
    pr, st = engine.pagerank(g, fixed_iterations=K)
    mark()
    ids_d = g.renumber_map()
    ids = torch.empty(ids_d.shape, dtype=ids_d.dtype, pin_memory=True); out = torch.empty(pr.shape, dtype=pr.dtype, pin_memory=True)
    ids.copy_(ids_d, non_blocking=True); out.copy_(pr, non_blocking=True)
    mark()
    del g, ds, dd, dv
    mark()
    names = ["h2d", "graph build", "pagerank call", "d2h", "free"]
    print(f"rep{rep}: " + "  ".join(f"{n} {1e3*(b-a):.1f} ms" for n, a, b in zip(names, t, t[1:])) +
          f"  | loop {st.elapsed_ms:.1f} ms  total {1e3*(t[-2]-t[0]):.1f} ms -> {ne*K/(t[-2]-t[0])/1e9:.1f} GTEPS", flush=True)
