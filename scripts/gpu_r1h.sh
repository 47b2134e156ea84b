#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/r1h
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_mg.py tests/test_gpu_banded.py -m gpu -x -q > $O/pytest_mg.log 2>&1; echo "pytest rc=$?"
tail -15 $O/pytest_mg.log
echo "elapsed $SECONDS"
