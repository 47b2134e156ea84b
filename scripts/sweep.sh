mkdir -p gpurun_out
(timeout 300 python scripts/gpu_debug.py 2>&1 | grep -E "rmat|spmv|bench" | grep -v "^\[graph" | tail -16
for c in 0 1 2 3 4 5; do echo "== CFG $c"; B200_PULL_CFG=$c timeout 200 python scripts/gpu_perf.py --scale 24 --reps 2 | tail -1; done
echo "== f64 s24"; timeout 200 python scripts/gpu_perf.py --scale 24 --dtype f64 --reps 2 | tail -1
echo "== f64 s26"; timeout 200 python scripts/gpu_perf.py --scale 26 --dtype f64 --reps 2 --iters 10 | tail -1
echo "== f32 s24 weighted"; timeout 200 python scripts/gpu_perf.py --scale 24 --weighted --reps 2 | tail -1
) 2>&1 | tee gpurun_out/perf4.log | tail -70
