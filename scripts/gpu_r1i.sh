#!/bin/bash
export PYTHONUNBUFFERED=1
O=gpurun_out/r1i
mkdir -p $O
N=${1:-2}
timeout 600 python -m pytest tests/test_gpu_mg.py -m gpu -x -q -k "torchrun" > $O/pytest_mg_n$N.log 2>&1; echo "pytest rc=$?"
tail -6 $O/pytest_mg_n$N.log
for b in 1 0; do
  B200_BANDED=$b timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
    bench.py --gpus $N --steps 100 --warmup 5 --no-e2e --no-cpu > $O/bench_n${N}_banded$b.json 2> $O/bench_n${N}_banded$b.err; echo "rc=$?"
  cat $O/bench_n${N}_banded$b.json
done
B200_BANDED=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 \
    bench.py --gpus $N --scale 26 --dtype f64 --steps 30 --warmup 3 --no-e2e --no-cpu > $O/bench_n${N}_s26_f64.json 2> $O/bench_n${N}_s26_f64.err; echo "rc=$?"
cat $O/bench_n${N}_s26_f64.json
echo "elapsed $SECONDS"
