#!/bin/bash
# One GPU call: parity of the banded path, sanitizer, bench of both paths, ncu evidence.
export PYTHONUNBUFFERED=1
O=gpurun_out/r1c
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/gpu.txt
echo "== banded parity tests"; date +%s
timeout 700 python -m pytest tests/test_gpu_banded.py -m gpu -x -q > $O/pytest_banded.log 2>&1; echo "banded tests rc=$?"
tail -5 $O/pytest_banded.log
echo "== sanitizer"; date +%s
B200_BANDED=1 B200_BAND_HOT_MIN=1 timeout 400 compute-sanitizer --tool memcheck --error-exitcode 9 \
  python scripts/gpu_banded_small.py 14 > $O/sanitizer.log 2>&1; echo "sanitizer rc=$?"
grep -E "ERROR SUMMARY|SMALL_OK|Invalid|spmv_err" $O/sanitizer.log | head -20
echo "== bench tiled"; date +%s
B200_BANDED=0 timeout 400 python bench.py --steps 100 --warmup 5 > $O/bench_tiled.json 2> $O/bench_tiled.err; echo "rc=$?"; cat $O/bench_tiled.json
echo "== bench banded"; date +%s
B200_BANDED=1 B200_BAND_VERBOSE=1 timeout 400 python bench.py --steps 100 --warmup 5 > $O/bench_banded.json 2> $O/bench_banded.err; echo "rc=$?"; cat $O/bench_banded.json; grep "b200" $O/bench_banded.err | head
echo "== bench banded f64 / s20"; date +%s
if [ $SECONDS -lt 480 ]; then
B200_BANDED=1 B200_BAND_VERBOSE=1 timeout 300 python bench.py --dtype f64 --steps 50 --warmup 5 --no-e2e --no-cpu > $O/bench_banded_f64.json 2> $O/bench_banded_f64.err; cat $O/bench_banded_f64.json
B200_BANDED=0 timeout 300 python bench.py --dtype f64 --steps 50 --warmup 5 --no-e2e --no-cpu > $O/bench_tiled_f64.json 2> $O/bench_tiled_f64.err; cat $O/bench_tiled_f64.json
B200_BANDED=1 timeout 200 python bench.py --scale 20 --steps 100 --warmup 5 --no-e2e --no-cpu > $O/bench_banded_s20.json 2>/dev/null; cat $O/bench_banded_s20.json
fi
echo "== ncu launch list (banded)"; date +%s
if [ $SECONDS -lt 560 ]; then
B200_BANDED=1 timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file $O/launches_banded.csv python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > $O/ncu_launch.log 2>&1; echo "rc=$?"
fi
echo "== ncu full (banded, one launch of each kernel)"; date +%s
if [ $SECONDS -lt 640 ]; then
B200_BANDED=1 timeout 500 ncu --set full --clock-control none --import-source on \
  -k regex:'band_reduce_kernel|pull_tiles_kernel' --launch-skip 6 -c 2 -o $O/banded_full -f \
  python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > $O/ncu_full.log 2>&1; echo "rc=$?"
fi
ls -la $O; date +%s; echo "elapsed $SECONDS"
