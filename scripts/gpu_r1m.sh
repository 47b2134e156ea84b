#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/r1m
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_banded.py -m gpu -x -q 2>&1 | tail -3
B200_BAND_KERNEL=2 timeout 200 python scripts/gpu_quick.py 24 k2p 2>&1 | tee $O/quick_k2p.log
B200_BAND_KERNEL=1 timeout 200 python scripts/gpu_quick.py 24 k1-w4096 2>&1 | tee $O/quick_k1.log
echo "elapsed $SECONDS"
