#!/bin/bash
# 1-GPU evidence batch: precision trajectory, GEMM A/B of the TMA-store epilogue, wide 1-GPU step, bench line.
mkdir -p gpurun_out
timeout 200 python scripts/precision_trajectory.py --out gpurun_out/r2_precision_trajectory.json 2>&1 | tail -3
for e in 0 1; do
  DMLP_GEMM_EPI2=$e timeout 150 python scripts/bench_gemm_large.py 8192x784x8192 > gpurun_out/r2_gemm_k784_epi2_$e.jsonl 2>&1
  echo "EPI2=$e"; cut -c1-260 gpurun_out/r2_gemm_k784_epi2_$e.jsonl
done
timeout 200 python bench.py --model wide --steps 30 --warmup 5 > gpurun_out/r2_bench_wide_1gpu.json 2> gpurun_out/r2_bench_wide_1gpu.err
echo "wide rc=$?"; python -c "import json; d=json.load(open('gpurun_out/r2_bench_wide_1gpu.json')); print('wide ms/step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['verified'])"
timeout 200 python bench.py --steps 2000 --warmup 200 > gpurun_out/r2_bench_small_1gpu.json 2> gpurun_out/r2_bench_small_1gpu.err
echo "small rc=$?"; python -c "import json; d=json.load(open('gpurun_out/r2_bench_small_1gpu.json')); print('small us/step', d['ms_per_step']*1e3, 'e2e', d['e2e']['ms_per_step']*1e3, d['verified'])"
