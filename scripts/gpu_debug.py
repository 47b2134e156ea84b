"""Development aid: run a battery of parity checks on the GPU and print diagnostics."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from cugraph_forked_b200 import engine
from oracle import cpu_oracle, graph_oracle
from tests import util

cpu_oracle.build()
h = engine.Handle()


def check_graph(name, src, dst, w, vertices=None):
    t0 = time.time()
    g = engine.Graph(h, torch.from_numpy(src), torch.from_numpy(dst),
                     None if w is None else torch.from_numpy(w),
                     vertices=None if vertices is None else torch.from_numpy(vertices))
    torch.cuda.synchronize()
    t1 = time.time()
    labels, seg, nmaj, nmin = graph_oracle.renumber_sg(dst, src, vertices)
    offs, idx, ww = graph_oracle.compress_sorted(nmaj, nmin, w, 0, len(labels))
    ok_map = np.array_equal(g.renumber_map().cpu().numpy(), labels)
    ok_seg = g.segment_offsets() == seg
    ok_off = np.array_equal(g.offsets().cpu().numpy().astype(np.int64), offs)
    ok_idx = np.array_equal(g.indices().cpu().numpy(), idx)
    ok_w = True if w is None else np.array_equal(g.weights().cpu().numpy(), ww)
    print(f"[graph {name}] V={len(labels)} E={len(src)} build={t1-t0:.3f}s map={ok_map} seg={ok_seg} "
          f"off={ok_off} idx={ok_idx} w={ok_w} seg={g.segment_offsets()}", flush=True)
    return g, labels


def check_pagerank(name, g, labels, src, dst, w, nv, dtype=np.float32, **kw):
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    t0 = time.time()
    pr, st = engine.pagerank(g, dtype=tdt, **kw)
    torch.cuda.synchronize()
    t1 = time.time()
    ref64, it64 = util.oracle_pagerank(src, dst, w, nv, dtype=np.float64,
                                       fixed_iterations=st.iterations)
    ext = util.to_external(pr.cpu().numpy(), labels, nv)
    mask = ~np.isnan(ext)
    err = util.rel_err(ext[mask], ref64[mask])
    top_ok = np.array_equal(util.top_k(np.where(mask, ext, -1)), util.top_k(np.where(mask, ref64, -1)))
    ref, it = util.oracle_pagerank(src, dst, w, nv, dtype=dtype, epsilon=kw.get("epsilon", 1e-6),
                                   alpha=kw.get("alpha", 0.85))
    print(f"[pagerank {name} {np.dtype(dtype).name}] iters gpu={st.iterations} oracle={it} conv={st.converged} "
          f"diff={st.last_diff:.3e} max_rel_err_vs_f64@same_iters={err.max():.3e} top100={top_ok} "
          f"sum={np.nansum(ext):.7f} loop_ms={st.elapsed_ms:.3f} wall={t1-t0:.3f}s", flush=True)


for name in ["karate.mtx", "netscience.mtx", "dolphins.mtx"]:
    src, dst, w, nv, sym = util.load_mtx(name)
    for weighted in (False, True):
        ww = w if weighted else None
        g, labels = check_graph(name + ("/w" if weighted else ""), src, dst, ww)
        check_pagerank(name, g, labels, src, dst, ww, nv, np.float32)
        check_pagerank(name, g, labels, src, dst, ww, nv, np.float64)

for scale in (10, 14, 18):
    src, dst = util.rmat(scale)
    nv = 1 << scale
    verts = np.arange(nv, dtype=np.int32)
    gs, gd = engine.generate_rmat_edgelist(h, scale, len(src))
    print(f"[rmat s{scale}] generator bit-exact: {np.array_equal(gs.cpu().numpy(), src) and np.array_equal(gd.cpu().numpy(), dst)}")
    g, labels = check_graph(f"rmat{scale}", src, dst, None, verts)
    check_pagerank(f"rmat{scale}", g, labels, src, dst, None, nv, np.float32)
    check_pagerank(f"rmat{scale}", g, labels, src, dst, None, nv, np.float64)
    x = torch.rand(nv, device="cuda")
    y = engine.spmv_pull(g, x, 2.0, 0.5).cpu().numpy().astype(np.float64)
    offs = g.offsets().cpu().numpy(); idx = g.indices().cpu().numpy()
    xs = x.cpu().numpy().astype(np.float64)
    yref = 0.5 + 2.0 * np.add.reduceat(np.concatenate([xs[idx], [0.0]]), np.minimum(offs[:-1], len(idx)))
    yref[np.diff(offs) == 0] = 0.5
    print(f"[spmv s{scale}] max rel err {util.rel_err(y, yref).max():.3e}")

# benchmark-ish: s20 fixed iterations
for scale in (20,):
    ne = 16 << scale
    gs, gd = engine.generate_rmat_edgelist(h, scale, ne)
    verts = torch.arange(1 << scale, dtype=torch.int32, device="cuda")
    t0 = time.time(); g = engine.Graph(h, gs, gd, vertices=verts); torch.cuda.synchronize()
    print(f"[bench s{scale}] build {time.time()-t0:.3f}s seg={g.segment_offsets()}")
    for _ in range(3):
        pr, st = engine.pagerank(g, fixed_iterations=20)
        print(f"[bench s{scale}] 20 iters {st.elapsed_ms:.3f} ms -> {ne*20/st.elapsed_ms/1e6:.1f} GTEPS launches={st.gpu_launches}")
    pr, st = engine.pagerank(g)
    print(f"[bench s{scale}] converge: iters={st.iterations} {st.elapsed_ms:.3f} ms -> {ne*st.iterations/st.elapsed_ms/1e6:.1f} GTEPS sum={pr.double().sum().item():.6f}")
