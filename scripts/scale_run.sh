#!/bin/bash
# bench.py at N = 1, 2, 4, 8 (as many as the box has) for the given --fused modes; one JSON line each under gpurun_out/.
MODES=${MODES:-"-1 4"}
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for n in 1 2 4 8; do
  [ $n -le $NG ] || continue
  for m in $MODES; do
    [ $n -eq 1 ] && [ "$m" != "-1" ] && continue
    out=gpurun_out/scale_n${n}_f${m}.json
    if [ $n -eq 1 ]; then
      timeout 300 python bench.py --gpus 1 --steps 2000 --warmup 200 --fused $m > $out 2> gpurun_out/scale_n${n}_f${m}.err
    else
      timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2970$n \
        bench.py --gpus $n --steps 2000 --warmup 200 --fused $m > $out 2> gpurun_out/scale_n${n}_f${m}.err
    fi
    echo "n=$n fused=$m rc=$? $(python -c "import json,sys; d=json.load(open('$out')); print('us/step', round(d['ms_per_step']*1e3,2), 'e2e', round(d['e2e']['ms_per_step']*1e3,2), 'value', round(d['value']))" 2>/dev/null)"
  done
done
