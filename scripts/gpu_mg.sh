#!/bin/bash
# multi-GPU: parity worker + bench lines with per-phase trace; usage: gpu_mg.sh <outdir> <N> [scale dtype]...
export PYTHONUNBUFFERED=1
O=gpurun_out/$1; N=$2; shift 2
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
[ -z "$SKIP_WORKER" ] && timeout 900 $TR tests/mg_worker.py --scale 12 > $O/mg_worker_n$N.log 2>&1; echo "mg_worker rc=$? ok=$(grep -c MG_WORKER_OK $O/mg_worker_n$N.log)"; grep -v MG_WORKER_OK $O/mg_worker_n$N.log | tail -5
run() { sc=$1; dt=$2; tag=$3; shift 3
  env "$@" B200_MG_TRACE=1 timeout 900 $TR bench.py --gpus $N --scale $sc --dtype $dt --steps 50 --warmup 5 --no-e2e --no-cpu > $O/bench_n${N}_s${sc}_${dt}_$tag.json 2> $O/bench_n${N}_s${sc}_${dt}_$tag.err; grep "mg trace" $O/bench_n${N}_s${sc}_${dt}_$tag.err | head -1; python -c "
import json
d=json.loads([l for l in open('$O/bench_n${N}_s${sc}_${dt}_$tag.json') if l.startswith('{')][-1]); print('N=$N s$sc $dt $tag: GTEPS', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'kernel_ms', round(d['roofline']['kernel_ms_avg'],4), 'parity', d['parity_check']['pass'], d['parity_check']['max_rel_err'])"
}
while [ $# -ge 2 ]; do
  run $1 $2 fused A=1
  run $1 $2 perpart B200_MG_FUSED=0
  shift 2
done
echo "elapsed $SECONDS"
