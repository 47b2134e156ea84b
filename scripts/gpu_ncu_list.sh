#!/bin/bash
# per-kernel device times (ncu launch list) of the SpMV kernels for "<scale> <dtype>" pairs
# usage: gpu_ncu_list.sh <outdir> "24 f32" "26 f64" ...
export PYTHONUNBUFFERED=1
O=gpurun_out/$1; shift
mkdir -p $O
for cfg in "$@"; do set -- $cfg
  python bench.py --scale $1 --dtype $2 --steps 3 --warmup 3 --no-e2e --no-cpu --no-parity > $O/plain_s$1_$2.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'band|pull_tiles|pull_tail|sell' -c 24 --csv --log-file $O/launches_s$1_$2.csv python bench.py --scale $1 --dtype $2 --steps 3 --warmup 3 --no-e2e --no-cpu --no-parity > $O/ncu_s$1_$2.log 2>&1; echo "ncu s$1 $2 rc=$?"
  grep -h "band\|pull_tiles\|pull_tail\|sell" $O/launches_s$1_$2.csv | awk -F'","' '{print substr($5,1,60), $NF}' | tail -6
done
echo "elapsed $SECONDS"
