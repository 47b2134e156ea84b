#!/bin/bash
# GPU call: banded v2 (work queue + lighter step body): parity, sweep, bench, ncu.
export PYTHONUNBUFFERED=1
O=gpurun_out/r1d
mkdir -p $O
echo "== banded parity tests"
timeout 600 python -m pytest tests/test_gpu_banded.py -m gpu -x -q > $O/pytest_banded.log 2>&1; echo "banded tests rc=$?"
tail -5 $O/pytest_banded.log
echo "== sweep f32"
timeout 500 python scripts/gpu_sweep_banded.py 24 f32 > $O/sweep_f32.log 2>&1; echo "rc=$?"; cat $O/sweep_f32.log
echo "== bench banded"
B200_BANDED=1 B200_BAND_VERBOSE=1 timeout 400 python bench.py --steps 100 --warmup 5 > $O/bench_banded.json 2> $O/bench_banded.err; echo "rc=$?"; cat $O/bench_banded.json; grep "b200" $O/bench_banded.err | head -3
B200_BANDED=1 timeout 300 python bench.py --dtype f64 --steps 50 --warmup 5 --no-e2e --no-cpu > $O/bench_banded_f64.json 2> $O/bench_banded_f64.err; cat $O/bench_banded_f64.json
if [ $SECONDS -lt 520 ]; then
echo "== ncu launch list (banded)"
B200_BANDED=1 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file $O/launches_banded.csv python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > $O/ncu_launch.log 2>&1; echo "rc=$?"
fi
if [ $SECONDS -lt 600 ]; then
echo "== ncu full (banded)"
B200_BANDED=1 timeout 400 ncu --set full --clock-control none --import-source on \
  -k regex:'band_reduce_kernel|pull_tiles_kernel' --launch-skip 6 -c 2 -o $O/banded_full -f \
  python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > $O/ncu_full.log 2>&1; echo "rc=$?"
fi
echo "elapsed $SECONDS"
