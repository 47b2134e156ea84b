#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/r1k
mkdir -p $O
echo "== parity, kernel 2 (default)"
timeout 600 python -m pytest tests/test_gpu_banded.py -m gpu -x -q 2>&1 | tail -3
echo "== timing"
B200_BAND_KERNEL=1 timeout 200 python scripts/gpu_quick.py 24 k1 2>&1 | tee $O/quick_k1.log
B200_BAND_KERNEL=2 timeout 200 python scripts/gpu_quick.py 24 k2 2>&1 | tee $O/quick_k2.log
for w in 3072 4096; do
B200_LIBRARY=$PWD/cugraph_forked_b200/variants/libcugraph_b200_w$w.so timeout 200 python scripts/gpu_quick.py 24 k2-w$w u32,w32 2>&1 | tee $O/quick_k2_w$w.log
done
echo "elapsed $SECONDS"
