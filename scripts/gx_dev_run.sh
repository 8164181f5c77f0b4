#!/bin/bash
# Development run of the global-batch data-parallel mode on N GPUs: oracle/soak/failure worker, then bench lines.
N=${1:-2}
mkdir -p gpurun_out
export DMLP_GX_SOAK_STEPS=${DMLP_GX_SOAK_STEPS:-3000}
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29631 \
  tests/mp_gx_worker.py > gpurun_out/gx_worker_${N}.log 2>&1
echo "gx worker rc=$?"; tail -n 25 gpurun_out/gx_worker_${N}.log
timeout 300 python bench.py --gpus 1 --steps 2000 --warmup 200 > gpurun_out/bench_1.json 2> gpurun_out/bench_1.err
echo "bench1 rc=$?"; cat gpurun_out/bench_1.json
for n in 2 4 8; do
  if [ $n -le $N ]; then
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2964$n \
      bench.py --gpus $n --steps 2000 --warmup 200 > gpurun_out/bench_$n.json 2> gpurun_out/bench_$n.err
    echo "bench$n rc=$?"; cat gpurun_out/bench_$n.json; tail -n 5 gpurun_out/bench_$n.err
  fi
done
