#!/bin/bash
# 1/2/4/8-GPU runs of bench.py (the driver does the same at round end)
for wl in u8 bin u8batch; do
for n in 8 4 2 1; do
  st=300; [ $wl = u8batch ] && st=10
  if [ $n = 1 ]; then python bench.py --gpus 1 --workload $wl --steps $st --no-cpu > gpurun_out/scale_${wl}_n$n.log 2>&1
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --workload $wl --steps $st --no-cpu > gpurun_out/scale_${wl}_n$n.log 2>&1; fi
  tail -1 gpurun_out/scale_${wl}_n$n.log | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$wl', d['n_gpus'], 'value', '%.4g'%d['value'], 'ms/step', round(d['ms_per_step'],4), 'e2e', '%.4g'%d['e2e']['value'])
except Exception as e: print('$wl $n FAILED', e)"
done; done
