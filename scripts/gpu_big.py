"""Development aid: scale-27 single-GPU run (2^31 edges -> 64-bit offsets) checked through
size-independent properties: sum(pr) = 1, offsets end at E, SpMV linearity, agreement of the
tiled kernel with a torch segmented sum on a sample of rows."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cugraph_forked_b200 import engine

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 27
nv, ne = 1 << scale, 16 << scale
h = engine.Handle()
t0 = time.time()
src, dst = engine.generate_rmat_edgelist(h, scale, ne)
verts = torch.arange(nv, dtype=torch.int32, device="cuda")
g = engine.Graph(h, src, dst, vertices=verts)
torch.cuda.synchronize()
print(f"s{scale}: V={nv} E={ne} gen+build {time.time()-t0:.2f}s edge_type={g.view().edge_type} seg={g.segment_offsets()}", flush=True)
del src, dst
off = g.offsets()
assert off.dtype == (torch.int64 if ne >= 2**31 - 1 else torch.int32), off.dtype
assert int(off[-1]) == ne and int(off[0]) == 0
assert bool((off[1:] >= off[:-1]).all())
pr, st = engine.pagerank(g, fixed_iterations=10)
s = float(pr.double().sum())
print(f"pagerank 10 iterations: {st.elapsed_ms/10:.3f} ms/iter {ne*10/st.elapsed_ms/1e6:.1f} GTEPS  sum={s:.9f}", flush=True)
assert abs(s - 1.0) < 1e-4
# raw SpMV against a segmented sum on sampled rows
x = torch.rand(nv, device="cuda", dtype=torch.float64)
y = engine.spmv_pull(g, x)
idx = g.indices()
rows = torch.cat([torch.arange(0, 64, device="cuda"), torch.randint(0, nv, (4096,), device="cuda")])
for r in rows.tolist()[:512]:
    b, e = int(off[r]), int(off[r + 1])
    want = float(x[idx[b:e].long()].sum())
    got = float(y[r])
    assert abs(got - want) <= 1e-9 * max(1.0, abs(want)), (r, got, want)
print("BIG_OK", flush=True)
