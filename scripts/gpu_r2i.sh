#!/bin/bash
# full GPU suite + default bench line (parity_check, cpu_baseline on the same graph) + reference arm
export PYTHONUNBUFFERED=1
O=gpurun_out/r2i
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; cat $O/bench.json; tail -3 $O/bench.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"; cat $O/bench_ref.json
timeout 600 python bench.py --scale 26 --dtype f64 --steps 20 --warmup 3 --no-cpu --no-e2e > $O/bench_s26_f64.json 2> $O/bench_s26.err; echo "s26 rc=$?"; cat $O/bench_s26_f64.json
echo "elapsed $SECONDS"
