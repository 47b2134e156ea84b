#!/bin/bash
# usage: gpu_mg2.sh <outdir> <N> <worker:0|1> [scale dtype tag env...;]   (one bench config per extra arg group via RUNS env)
export PYTHONUNBUFFERED=1
O=gpurun_out/$1; N=$2; W=$3
mkdir -p $O
TR="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
if [ "$W" = "1" ]; then $TR tests/mg_worker.py --scale 12 > $O/mg_worker_n$N.log 2>&1; echo "mg_worker rc=$? ok=$(grep -c MG_WORKER_OK $O/mg_worker_n$N.log)"; grep -i "error\|assert" $O/mg_worker_n$N.log | head -5; fi
run() { sc=$1; dt=$2; tag=$3; shift 3
  env B200_MG_TRACE=2 B200_BAND_VERBOSE=1 "$@" $TR bench.py --gpus $N --scale $sc --dtype $dt --steps 50 --warmup 5 --no-e2e --no-cpu > $O/bench_n${N}_s${sc}_${dt}_$tag.json 2> $O/bench_n${N}_s${sc}_${dt}_$tag.err; grep "mg trace\] rank" $O/bench_n${N}_s${sc}_${dt}_$tag.err | tail -$N | cut -c16-200; grep "sell layout" $O/bench_n${N}_s${sc}_${dt}_$tag.err | tail -1 | cut -c1-420; python -c "
import json
d=json.loads([l for l in open('$O/bench_n${N}_s${sc}_${dt}_$tag.json') if l.startswith('{')][-1]); print('N=$N s$sc $dt $tag: GTEPS', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'kernel_ms', round(d['roofline']['kernel_ms_avg'],4), 'parity', d['parity_check']['pass'], d['parity_check']['max_rel_err'])" || tail -5 $O/bench_n${N}_s${sc}_${dt}_$tag.err
}
eval "$RUNS"
echo "elapsed $SECONDS"
