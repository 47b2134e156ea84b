#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/n4
mkdir -p $O
B200_BANDED=1 B200_BAND_MIN_EDGES=512 B200_BAND_HOT_MIN=64 timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29604 tests/mg_worker.py --scale 12 > $O/mg_worker_n4.log 2>&1; echo "worker rc=$? ok=$(grep -c MG_WORKER_OK $O/mg_worker_n4.log)"
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29614 \
    bench.py --gpus 4 --steps 100 --warmup 5 --no-e2e --no-cpu > $O/bench_n4.json 2> $O/bench_n4.err; echo "rc=$?"
cat $O/bench_n4.json
echo "elapsed $SECONDS"
