#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/s26
mkdir -p $O
B200_BAND_VERBOSE=1 timeout 100 python bench.py --scale 26 --dtype f64 --steps 20 --warmup 3 --no-e2e --no-cpu > $O/bench_s26_f64_n1.json 2> $O/bench_s26_f64_n1.err; echo "rc=$?"
cat $O/bench_s26_f64_n1.json; grep b200 $O/bench_s26_f64_n1.err | head -2
echo "elapsed $SECONDS"
