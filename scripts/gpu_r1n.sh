#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/r1n
mkdir -p $O
timeout 100 python scripts/gpu_quick.py 24 g16 u32,u64 2>&1 | tee $O/quick_g16.log
for g in 12 20 24; do
B200_LIBRARY=$PWD/cugraph_forked_b200/variants/libcugraph_b200_g$g.so timeout 100 python scripts/gpu_quick.py 24 g$g u32,u64 2>&1 | tee $O/quick_g$g.log
done
echo "elapsed $SECONDS"
