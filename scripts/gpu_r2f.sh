#!/bin/bash
# sell layout v2: uniform phase-1 tasks, pipelined phase 2, slot-order switch
export PYTHONUNBUFFERED=1
O=gpurun_out/r2f
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_banded.py tests/test_gpu_centrality.py tests/test_gpu_configs.py -m gpu -x -q -k "not s24 and not scale27" > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
run() { # scale dtype tag [env...]
  sc=$1; dt=$2; tag=$3; shift 3
  env "$@" B200_BAND_VERBOSE=1 timeout 600 python bench.py --scale $sc --dtype $dt --steps 20 --warmup 3 --no-e2e --no-cpu > $O/bench_s${sc}_${dt}_$tag.json 2> $O/bench_s${sc}_${dt}_$tag.err; echo "bench s$sc $dt $tag rc=$?"; grep "layout" $O/bench_s${sc}_${dt}_$tag.err | head -1 | sed 's/.*slots=/slots=/'; python -c "
import json,sys
d=json.load(open('$O/bench_s${sc}_${dt}_$tag.json')); print('   GTEPS', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],3))"
}
run 24 f32 def A=1
run 24 f32 bm B200_SELL_SLOT_BYTES=1
run 24 f64 def A=1
run 24 f64 bm B200_SELL_SLOT_BYTES=1
run 26 f64 def A=1
run 26 f64 rm B200_SELL_SLOT_BYTES=100000000000
run 20 f32 def A=1
echo "elapsed $SECONDS"
