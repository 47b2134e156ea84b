#!/bin/bash
# evidence for the committed default: full GPU suite, smoke, bench lines, ncu launch list and
# full capture of the SpMV kernels
export PYTHONUNBUFFERED=1
O=gpurun_out/final
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 400 python bench.py > $O/bench.json 2> $O/bench.err; echo "rc=$?"; cat $O/bench.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file $O/launches.csv python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > $O/ncu_launch.log 2>&1; echo "rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on \
  -k regex:'band_reduce2_kernel|pull_tiles_kernel' --launch-skip 6 -c 2 -o $O/full -f \
  python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > $O/ncu_full.log 2>&1; echo "rc=$?"
timeout 200 python scripts/gpu_quick.py 24 final 2>&1 | tee $O/quick.log
echo "elapsed $SECONDS"
