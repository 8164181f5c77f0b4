#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
for mode in 5 4; do
DMLP_PROF_FUSED_MODE=$mode timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2965$mode \
  scripts/prof_fused_dp.py > gpurun_out/gx_prof_${N}_mode$mode.log 2>&1
echo "prof mode $mode rc=$?"; grep phase_us gpurun_out/gx_prof_${N}_mode$mode.log
done
timeout 100 python scripts/prof_fused_dp.py > gpurun_out/gx_prof_1.log 2>&1; grep phase_us gpurun_out/gx_prof_1.log
