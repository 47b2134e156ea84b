mkdir -p gpurun_out
(for c in 0 1 2 3 4 5; do for kb in 160 184; do echo "== CFG $c budget $kb"; B200_SMEM_BUDGET_KB=$kb B200_PULL_CFG=$c timeout 200 python scripts/gpu_perf.py --scale 24 --reps 2 | tail -1; done; done
) 2>&1 | tee gpurun_out/sweep3.log | tail -70
