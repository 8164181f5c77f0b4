#!/bin/bash
# One 8-GPU box visit: multi-GPU tests, headline bench at N=8 (default = persistent data-parallel launch, and the
# two-kernel chain for comparison), N=4, wide model at 8.
set -u
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_multigpu.py -m gpu -q -x 2>&1 | tail -5 | tee $O/mg_tests_8gpu.txt
timeout 120 $TR --nproc-per-node 8 --master-port 29601 bench.py --gpus 8 --steps 4000 --warmup 400 2>/dev/null | tee $O/bench_small_8gpu.json | cut -c1-200
timeout 120 $TR --nproc-per-node 8 --master-port 29607 bench.py --gpus 8 --steps 4000 --warmup 400 --fused 1 2>/dev/null | tee $O/bench_small_8gpu_two_kernel_chain.json | cut -c1-200
timeout 120 $TR --nproc-per-node 4 --master-port 29602 bench.py --gpus 4 --steps 4000 --warmup 400 2>/dev/null | tee $O/bench_small_4gpu.json | cut -c1-200
timeout 120 $TR --nproc-per-node 8 --master-port 29603 bench.py --gpus 8 --model wide --steps 30 --warmup 5 2>/dev/null | tee $O/bench_wide_8gpu.json | cut -c1-200
echo SUITE_DONE
