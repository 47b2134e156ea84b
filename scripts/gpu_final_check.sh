#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/final2; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --scale 20 --steps 20 --warmup 5 --no-cpu > $O/bench_s20_f32.json 2>/dev/null
python - <<'P'
import json,glob
for f in sorted(glob.glob("gpurun_out/final2/bench*.json")):
    d=json.loads([l for l in open(f) if l.startswith("{")][-1]); r=d["roofline"]
    print(f.split("/")[-1], "GTEPS", round(d["value"],2), "frac", round(r["frac"],3), "path", d["config"]["pull_path"], "e2e", d.get("e2e") and round(d["e2e"]["value"],1), "parity", d["parity_check"]["pass"])
P
echo "elapsed $SECONDS"
