#!/bin/bash
# Round 2, call c: where the time goes at the target config (s26 f64) and at s24 f64 / f32:
# layout statistics (B200_BAND_VERBOSE) + ncu launch lists (per-kernel device time).
export PYTHONUNBUFFERED=1
O=gpurun_out/r2c
mkdir -p $O
for cfg in "26 f64" "24 f64" "24 f32"; do set -- $cfg
  B200_BAND_VERBOSE=1 timeout 600 python bench.py --scale $1 --dtype $2 --steps 10 --warmup 3 --no-e2e --no-cpu > $O/bench_s$1_$2.json 2> $O/bench_s$1_$2.err; echo "bench s$1 $2 rc=$?"; grep banded $O/bench_s$1_$2.err | head -3
done
python bench.py --scale 26 --dtype f64 --steps 3 --warmup 3 --no-e2e --no-cpu > $O/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'band_reduce|pull_tiles|pull_tail' -c 18 --csv --log-file $O/launches_s26_f64.csv python bench.py --scale 26 --dtype f64 --steps 3 --warmup 3 --no-e2e --no-cpu > $O/ncu_s26.log 2>&1; echo "ncu s26 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'band_reduce|pull_tiles|pull_tail' -c 18 --csv --log-file $O/launches_s24_f64.csv python bench.py --scale 24 --dtype f64 --steps 3 --warmup 3 --no-e2e --no-cpu > $O/ncu_s24.log 2>&1; echo "ncu s24 rc=$?"
grep -h "band_reduce\|pull_tiles\|pull_tail" $O/launches_s26_f64.csv | awk -F'","' '{print $5, $NF}' | tail -9
grep -h "band_reduce\|pull_tiles\|pull_tail" $O/launches_s24_f64.csv | awk -F'","' '{print $5, $NF}' | tail -9
echo "elapsed $SECONDS"
