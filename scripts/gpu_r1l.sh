#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/r1l
mkdir -p $O
timeout 400 ncu --set full --clock-control none --import-source on \
  -k regex:'band_reduce2_kernel' --launch-skip 3 -c 1 -o $O/k2_full -f \
  python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > $O/ncu_full.log 2>&1; echo "rc=$?"
echo "elapsed $SECONDS"
