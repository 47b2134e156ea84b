#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/r1f
mkdir -p $O
echo "== f64 / weighted config check"
cat > /tmp/cfgcheck.py <<'PY'
import os, sys, torch
sys.path.insert(0, os.getcwd())
from cugraph_forked_b200 import engine
h = engine.Handle()
scale = 24
nv, ne = 1 << scale, 16 << scale
s, d = engine.generate_rmat_edgelist(h, scale, ne, seed=0)
verts = torch.arange(nv, dtype=torch.int32, device="cuda")
for weighted in (False, True):
    w = torch.rand(ne, device="cuda", dtype=torch.float32) if weighted else None
    for dt in (torch.float32, torch.float64):
        for bc in (0, 1):
            os.environ["B200_BANDED_CFG"] = str(bc)
            g = engine.Graph(h, s, d, w, vertices=verts)
            engine.pagerank(g, dtype=dt, fixed_iterations=5)
            _, st = engine.pagerank(g, dtype=dt, fixed_iterations=30)
            print(f"weighted={weighted} dtype={dt} banded_cfg={bc}: {st.elapsed_ms/30:.4f} ms/iter", flush=True)
            del g
PY
timeout 300 python /tmp/cfgcheck.py > $O/cfgcheck.log 2>&1; echo "rc=$?"; cat $O/cfgcheck.log
echo "== full GPU suite (banded default on)"
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -8 $O/pytest_gpu.log
echo "== bench default"
timeout 400 python bench.py > $O/bench.json 2> $O/bench.err; echo "rc=$?"; cat $O/bench.json
echo "elapsed $SECONDS"
