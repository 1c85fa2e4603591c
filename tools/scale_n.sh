#!/bin/bash
# bench.py on N GPUs (N = $1) for the single-query workloads with the default exchange
n=${1:-8}
for wl in u8 bin pq; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29900+n)) bench.py --gpus $n --workload $wl --steps 500 --no-cpu > gpurun_out/scale2_${wl}_n$n.log 2>&1
  tail -1 gpurun_out/scale2_${wl}_n$n.log | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$wl', d['n_gpus'], 'value %.4g'%d['value'], 'ms/step', round(d['ms_per_step'],4), 'e2e ms', round(d['e2e']['ms_per_step'],4))
except Exception as e: print('$wl FAILED', e)"
done
