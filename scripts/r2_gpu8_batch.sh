#!/bin/bash
# 8-GPU batch: multi-GPU test file (8 ranks), scaling run, wide model with a failure drill, all-reduce sweep, NCCL-graph baseline.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
timeout 600 python -m pytest tests/test_gpu_multigpu.py -m gpu -x -q 2>&1 | tail -4
MODES="-1" timeout 400 bash scripts/scale_run.sh
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29711 \
  bench.py --gpus $NG --steps 2000 --warmup 200 --fault dead-rank:5@step:40 --watchdog-ms 1000 > gpurun_out/r2_bench_small_${NG}gpu_drill.json 2> gpurun_out/r2_bench_small_${NG}gpu_drill.err
echo "small drill rc=$?"; python -c "import json; d=json.load(open('gpurun_out/r2_bench_small_${NG}gpu_drill.json')); print(d['ms_per_step'], d['verified'], d['failure_drill'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29712 \
  bench.py --model wide --gpus $NG --steps 30 --warmup 5 --fault dead-rank:5@step:10 --watchdog-ms 2000 > gpurun_out/r2_bench_wide_${NG}gpu_drill.json 2> gpurun_out/r2_bench_wide_${NG}gpu_drill.err
echo "wide drill rc=$?"; python -c "import json; d=json.load(open('gpurun_out/r2_bench_wide_${NG}gpu_drill.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['verified'], d['failure_drill'])"; tail -n 3 gpurun_out/r2_bench_wide_${NG}gpu_drill.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29520 \
  scripts/bench_allreduce.py --max-bytes 1073741824 --out gpurun_out/r2_allreduce_sweep_${NG}gpu_graph.jsonl > gpurun_out/r2_sweep${NG}.log 2>&1
echo "sweep rc=$?"; tail -n 2 gpurun_out/r2_sweep${NG}.log | cut -c1-200
timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29531 \
  baseline/nccl_graph_mlp.py --steps 2000 --time-limit 60 > gpurun_out/r2_baseline_nccl_graph_${NG}gpu.json 2> /dev/null
echo "nccl graph rc=$?"; tail -n 1 gpurun_out/r2_baseline_nccl_graph_${NG}gpu.json | cut -c1-400
