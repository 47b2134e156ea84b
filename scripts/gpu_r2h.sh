#!/bin/bash
# L2 policy hints / fetch granularity / occupancy variants of phase 2
export PYTHONUNBUFFERED=1
O=gpurun_out/r2h
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_banded.py tests/test_gpu_centrality.py -m gpu -x -q -k "not scale27" > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
run() { # scale dtype tag [env...]
  sc=$1; dt=$2; tag=$3; shift 3
  env "$@" timeout 600 python bench.py --scale $sc --dtype $dt --steps 20 --warmup 3 --no-e2e --no-cpu > $O/bench_s${sc}_${dt}_$tag.json 2> $O/bench_s${sc}_${dt}_$tag.err; python -c "
import json,sys
d=json.load(open('$O/bench_s${sc}_${dt}_$tag.json')); print('s$sc $dt $tag: GTEPS', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],3))"
}
V=$PWD/cugraph_forked_b200/variants
for cfg in "26 f64" "24 f32"; do set -- $cfg
run $1 $2 def A=1
run $1 $2 fetch64 B200_L2_FETCH=64
run $1 $2 hot0 B200_HOT_MB=0
run $1 $2 hot20 B200_HOT_MB=20
run $1 $2 hot80 B200_HOT_MB=80
run $1 $2 c5 B200_LIBRARY=$V/libcugraph_b200_c5.so
run $1 $2 c4 B200_LIBRARY=$V/libcugraph_b200_c4.so
run $1 $2 c5x10 B200_LIBRARY=$V/libcugraph_b200_c5.so B200_SELL_CTAS_PER_SM=5
done
run 24 f64 def A=1
run 20 f32 def A=1
echo "elapsed $SECONDS"
