#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/r1g
mkdir -p $O
echo "== quick: default lib (kMinCell 3), banded cfg 1 / 0, tiled"
timeout 200 python scripts/gpu_quick.py 24 mc3 2>&1 | tee $O/quick_mc3.log
B200_BANDED_CFG=0 timeout 200 python scripts/gpu_quick.py 24 mc3-oldcfg u64,w32,w64 2>&1 | tee $O/quick_mc3_oldcfg.log
B200_BANDED=0 timeout 200 python scripts/gpu_quick.py 24 tiled u64,w32,w64 2>&1 | tee $O/quick_tiled.log
echo "== quick: kMinCell 2 variant"
B200_LIBRARY=$PWD/cugraph_forked_b200/variants/libcugraph_b200_mc2.so B200_BAND_VERBOSE=1 timeout 200 python scripts/gpu_quick.py 24 mc2 2>&1 | tee $O/quick_mc2.log
echo "== parity of the kMinCell 2 variant"
B200_LIBRARY=$PWD/cugraph_forked_b200/variants/libcugraph_b200_mc2.so timeout 400 python -m pytest tests/test_gpu_banded.py -m gpu -x -q 2>&1 | tail -3
echo "== full GPU suite (banded default on)"
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -8 $O/pytest_gpu.log
echo "elapsed $SECONDS"
