#!/bin/bash
# ncu launch list + one full capture of the SpMV kernels; usage: gpu_ncu_full.sh <outdir> <scale> <dtype>
export PYTHONUNBUFFERED=1
O=gpurun_out/$1; sc=$2; dt=$3
mkdir -p $O
python bench.py --scale $sc --dtype $dt --steps 3 --warmup 3 --no-e2e --no-cpu --no-parity > $O/plain_s${sc}_$dt.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'band|pull_tiles|pull_tail|sell' -c 24 --csv --log-file $O/launches_s${sc}_$dt.csv python bench.py --scale $sc --dtype $dt --steps 3 --warmup 3 --no-e2e --no-cpu --no-parity > $O/ncu_list.log 2>&1; echo "ncu list rc=$?"
grep -h "band\|pull_tiles\|pull_tail\|sell" $O/launches_s${sc}_$dt.csv | awk -F'","' '{print substr($5,1,60), $NF}' | tail -3
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'sell_band|sell_rows' -s 6 -c 2 -o $O/prof_s${sc}_$dt python bench.py --scale $sc --dtype $dt --steps 3 --warmup 3 --no-e2e --no-cpu --no-parity > $O/ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la $O
echo "elapsed $SECONDS"
