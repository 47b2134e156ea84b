#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/r1e
mkdir -p $O
echo "== sweep f32"
timeout 500 python scripts/gpu_sweep_banded.py 24 f32 > $O/sweep_f32.log 2>&1; echo "rc=$?"; cat $O/sweep_f32.log
echo "== sweep f64"
timeout 300 python scripts/gpu_sweep_banded.py 24 f64 > $O/sweep_f64.log 2>&1; echo "rc=$?"; cat $O/sweep_f64.log
echo "elapsed $SECONDS"
