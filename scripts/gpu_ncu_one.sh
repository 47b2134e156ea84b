#!/bin/bash
# one ncu --set full capture: gpu_ncu_one.sh <outdir> <scale> <dtype> <kernel regex> <skip>
export PYTHONUNBUFFERED=1
O=gpurun_out/$1; mkdir -p $O
CMD="python bench.py --scale $2 --dtype $3 --steps 3 --warmup 3 --no-e2e --no-cpu --no-parity"
$CMD > $O/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$4 -s $5 -c 1 -o $O/prof_s$2_$3_$4 $CMD > $O/ncu_full.log 2>&1; echo "ncu rc=$?"; ls -la $O | tail -3
