#!/bin/bash
# Round 2, first GPU call: (1) the full GPU suite - it includes the tests written after round 1's
# GPU budget was spent (tests/test_zz_gpu_python_callers.py, the hits_file_golden client);
# (the device text parser's first run is tests/test_zzz_gpu_ingest_device.py); (2) smoke + the default bench line; (3) the overlap micro-benchmark that decides the next
# kernel step; (4) host-side ingest timing on the box's cores.
export PYTHONUNBUFFERED=1
O=gpurun_out/r2a
mkdir -p $O
B200_TEST_DEVICE_INGEST=1 timeout 1200 python -m pytest tests -m gpu -q --runxfail > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 400 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; cat $O/bench.json
timeout 120 scripts/microbench/overlap_bench > $O/overlap.log 2>&1; echo "overlap rc=$?"; cat $O/overlap.log
B200_INGEST_VERBOSE=1 timeout 300 python scripts/ingest_bench.py 16000000 1 8 0 > $O/ingest.log 2>&1; echo "ingest rc=$?"; cat $O/ingest.log
echo "elapsed $SECONDS"
