#!/bin/bash
# round-end evidence on one GPU: full GPU suite, smoke, default bench line, reference arm, other configs
export PYTHONUNBUFFERED=1
O=gpurun_out/final1; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -E "pagerank_reference|max rel err" $O/pytest_gpu.log | tail -6; tail -3 $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
for cfg in "24 f64" "26 f64" "22 f32" "20 f32"; do set -- $cfg
  timeout 600 python bench.py --scale $1 --dtype $2 --steps 20 --warmup 5 --no-cpu --no-e2e > $O/bench_s$1_$2.json 2> $O/bench_s$1_$2.err
done
timeout 600 python bench.py --steps 100 --warmup 5 --no-cpu > $O/bench_steps100.json 2>/dev/null
python - <<'P'
import json,glob
for f in sorted(glob.glob("gpurun_out/final1/bench*.json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1])
        r=d.get("roofline") or {}
        print(f.split("/")[-1], "GTEPS", round(d["value"],2), "ms", None if d["ms_per_step"] is None else round(d["ms_per_step"],4), "frac", r.get("frac") and round(r["frac"],3), "e2e", d.get("e2e") and round(d["e2e"]["value"],1), "parity", (d.get("parity_check") or {}).get("pass"), "traffic", r.get("traffic"))
    except Exception as e: print(f, "ERR", e)
P
echo "elapsed $SECONDS"
