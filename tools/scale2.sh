#!/bin/bash
# N-GPU bench of the two exchange paths (N = $1)
n=${1:-2}
for wl in u8 bin pq; do for ex in peer nccl; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800+n)) bench.py --gpus $n --workload $wl --steps 500 --no-cpu --exchange $ex > gpurun_out/x_${wl}_${ex}_n$n.log 2>&1
  tail -1 gpurun_out/x_${wl}_${ex}_n$n.log | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$wl $ex', d['n_gpus'], 'value %.4g'%d['value'], 'ms/step', round(d['ms_per_step'],4), 'e2e ms', round(d['e2e']['ms_per_step'],4))
except Exception as e: print('$wl $ex FAILED', e)"
done; done
