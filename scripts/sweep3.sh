mkdir -p gpurun_out
(echo "== base"; timeout 200 python scripts/gpu_perf.py --scale 24 --reps 2 | tail -1
for v in hot64k hot256k hot1m prof; do echo "== $v"; B200_LIBRARY=$PWD/cugraph_forked_b200/variants/libcugraph_b200_$v.so timeout 200 python scripts/gpu_perf.py --scale 24 --reps 2 | tail -8 | grep -v "rep0\|^s24"; done
) 2>&1 | tee gpurun_out/sweep4.log | tail -70
