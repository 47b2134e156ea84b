N=$1
run() { echo "== $*"; env "$@" B200_MG_TRACE=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --steps 50 --warmup 5 --no-cpu --no-e2e 2>&1 | grep -E "^\{|50 iterations" | cut -c1-230; }
run A=1
run NCCL_NCHANNELS_PER_PEER=8
run NCCL_NCHANNELS_PER_PEER=16
run NCCL_NCHANNELS_PER_PEER=32 NCCL_MIN_P2P_NCHANNELS=16
run NCCL_NCHANNELS_PER_PEER=16 B200_MG_RESERVED_SMS=24
