#!/bin/bash
# Round 2, call d: first run of the sell layout (sell.cuh): parity tests, then bench lines.
export PYTHONUNBUFFERED=1
O=gpurun_out/r2d
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_banded.py tests/test_gpu_centrality.py tests/test_gpu_configs.py -m gpu -x -q -k "not s24 and not scale27" > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest.log
for cfg in "24 f32" "24 f64" "26 f64" "20 f32"; do set -- $cfg
  B200_BAND_VERBOSE=1 timeout 600 python bench.py --scale $1 --dtype $2 --steps 20 --warmup 3 --no-e2e --no-cpu > $O/bench_s$1_$2.json 2> $O/bench_s$1_$2.err; echo "bench s$1 $2 rc=$?"; grep "layout" $O/bench_s$1_$2.err | head -3; python -c "
import json,sys
d=json.load(open('$O/bench_s$1_$2.json')); print('GTEPS', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],3))"
done
echo "elapsed $SECONDS"
