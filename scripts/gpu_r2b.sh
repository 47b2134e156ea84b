#!/bin/bash
# Round 2, call b: the no-alloc entry point alone in a fresh process, then the whole GPU suite
# (new: tests/test_gpu_configs.py on the BASELINE configs, tests/test_zzzz_no_alloc.py last).
export PYTHONUNBUFFERED=1
O=gpurun_out/r2b
mkdir -p $O
timeout 600 python -m pytest tests/test_zzzz_no_alloc.py -m gpu -q > $O/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -5 $O/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q -s --durations=8 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -E "pagerank_reference|passed|failed|Error" $O/pytest_gpu.log | tail -12; tail -14 $O/pytest_gpu.log
echo "elapsed $SECONDS"
