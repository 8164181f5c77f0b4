#!/bin/bash
# 2-GPU development run of the wgrad reduce-scatter pipeline: worker test, then the wide model with and without it.
NG=${NG:-2}
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$NG --master-addr 127.0.0.1 --master-port 29623 tests/mp_rs_worker.py > gpurun_out/rs_worker.log 2>&1
echo "worker rc=$?"; grep -E "ok|OK|Error|error|assert" gpurun_out/rs_worker.log | head -20; tail -n 5 gpurun_out/rs_worker.log
for RS in ${RS_LIST:-1 0}; do
  DMLP_WGRAD_RS=$RS timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$NG --master-addr 127.0.0.1 --master-port 2963$RS \
    bench.py --model wide --gpus $NG --steps 20 --warmup 5 > gpurun_out/rs_bench_wide_${NG}gpu_rs$RS.json 2> gpurun_out/rs_bench_wide_${NG}gpu_rs$RS.err
  echo "bench rs=$RS rc=$?"; python -c "import json; d=json.load(open('gpurun_out/rs_bench_wide_${NG}gpu_rs$RS.json')); print(d['ms_per_step'], d['e2e'], d.get('verified'))"; tail -n 2 gpurun_out/rs_bench_wide_${NG}gpu_rs$RS.err
done
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$NG --master-addr 127.0.0.1 --master-port 29640 scripts/prof_wgrad_rs.py 2>&1 | grep -E "^----|^consumer|^  13|^   0|^   6" | head -14
