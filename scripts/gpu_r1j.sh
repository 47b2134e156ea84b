#!/bin/bash
# evidence for the committed default: bench line, ncu launch list, full capture of both kernels
export PYTHONUNBUFFERED=1
O=gpurun_out/r1j
mkdir -p $O
timeout 400 python bench.py > $O/bench.json 2> $O/bench.err; echo "rc=$?"; cat $O/bench.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file $O/launches.csv python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > $O/ncu_launch.log 2>&1; echo "rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on \
  -k regex:'band_reduce_kernel|pull_tiles_kernel' --launch-skip 6 -c 2 -o $O/full -f \
  python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > $O/ncu_full.log 2>&1; echo "rc=$?"
timeout 300 python bench.py --dtype f64 --steps 50 --no-e2e --no-cpu > $O/bench_f64.json 2>/dev/null; cat $O/bench_f64.json
timeout 300 python bench.py --scale 20 --no-e2e --no-cpu > $O/bench_s20.json 2>/dev/null; cat $O/bench_s20.json
timeout 300 python bench.py --impl reference --steps 1 --warmup 0 > $O/bench_reference.json 2>/dev/null; cat $O/bench_reference.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
echo "elapsed $SECONDS"
