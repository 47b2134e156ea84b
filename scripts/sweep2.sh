mkdir -p gpurun_out
(python -m pytest tests/test_gpu_centrality.py tests/test_gpu_graph.py -m gpu -x -q 2>&1 | tail -3
echo "== base"; timeout 200 python scripts/gpu_perf.py --scale 24 --reps 2 | tail -1
for v in noreduce ngnr g12; do echo "== $v"; B200_LIBRARY=$PWD/cugraph_forked_b200/variants/libcugraph_b200_$v.so timeout 200 python scripts/gpu_perf.py --scale 24 --reps 2 | tail -1; done
for c in 1 2 3; do echo "== base CFG $c"; B200_PULL_CFG=$c timeout 200 python scripts/gpu_perf.py --scale 24 --reps 2 | tail -1; done
echo "== f64 s24"; timeout 200 python scripts/gpu_perf.py --scale 24 --dtype f64 --reps 2 | tail -1
echo "== f32 s24 weighted"; timeout 200 python scripts/gpu_perf.py --scale 24 --weighted --reps 2 | tail -1
echo "== f64 s26"; timeout 200 python scripts/gpu_perf.py --scale 26 --dtype f64 --reps 2 --iters 10 | tail -1
) 2>&1 | tee gpurun_out/sweep2.log | tail -70
