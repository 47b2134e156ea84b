#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/hits
mkdir -p $O
timeout 90 python -m pytest tests/test_zz_gpu_hits.py tests/test_gpu_c_api.py -m gpu -q > $O/pytest_hits.log 2>&1; echo "pytest rc=$?"
tail -30 $O/pytest_hits.log
echo "elapsed $SECONDS"
