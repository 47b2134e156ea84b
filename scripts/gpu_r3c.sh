#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/r3c; mkdir -p $O
timeout 900 python -m pytest tests/test_zz_gpu_hits.py tests/test_gpu_c_api.py tests/test_gpu_centrality.py tests/test_gpu_banded.py tests/test_zzzz_no_alloc.py -m gpu -x -q -k "not scale27" > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest.log
python scripts/gpu_hits_bench.py 22 f32; python scripts/gpu_hits_bench.py 24 f32; python scripts/gpu_hits_bench.py 24 f64
echo "elapsed $SECONDS"
