#!/bin/bash
# 2-GPU batch: full GPU test-suite, graph-captured NCCL baseline, all-reduce sweep with every implementation.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
  baseline/nccl_graph_mlp.py --steps 2000 --time-limit 60 > gpurun_out/r2_baseline_nccl_graph_2gpu.json 2> gpurun_out/r2_baseline_nccl_graph_2gpu.err
echo "nccl graph 2gpu rc=$?"; tail -n 1 gpurun_out/r2_baseline_nccl_graph_2gpu.json | cut -c1-400
timeout 60 python baseline/nccl_graph_mlp.py --steps 2000 --time-limit 50 > gpurun_out/r2_baseline_nccl_graph_1gpu.json 2>/dev/null; tail -n 1 gpurun_out/r2_baseline_nccl_graph_1gpu.json | cut -c1-300
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29520 \
  scripts/bench_allreduce.py --max-bytes 1073741824 --out gpurun_out/r2_allreduce_sweep_2gpu_graph.jsonl > gpurun_out/r2_sweep2.log 2>&1
echo "sweep rc=$?"; tail -n 3 gpurun_out/r2_sweep2.log | cut -c1-200
