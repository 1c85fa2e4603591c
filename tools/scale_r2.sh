#!/bin/bash
# round 2: bench.py on N GPUs (N = $1) — u8 (C2) and binary (C3), all-to-all peer exchange and rooted gather, each line
# carrying the weak record, the strong record (BASELINE corpus split over N) and the global top-k check; then
# tools/peer_check.py at world N.  Lines are appended to gpurun_out/r2_strong_<wl>.jsonl.
n=${1:-8}
port=$((29800 + n))
DRY=${DRY:-}
for wl in u8 bin; do
  for ex in peer gather; do
    if [ "$n" = 1 ]; then
      [ "$ex" = gather ] && continue
      cmd="python bench.py --workload $wl --no-cpu --no-others --steps 1000"
    else
      cmd="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --workload $wl --exchange $ex --no-cpu --no-others --steps 1000"
      port=$((port + 1))
    fi
    if [ -n "$DRY" ]; then echo "$cmd"; continue; fi
    timeout 400 $cmd > gpurun_out/tmp_line.json 2> gpurun_out/scale_r2_${wl}_${ex}_n$n.err
    tail -1 gpurun_out/tmp_line.json >> gpurun_out/r2_strong_${wl}_n${n}.jsonl
    tail -1 gpurun_out/tmp_line.json | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); s=d.get('strong') or {}
    print('$wl $ex n', d['n_gpus'], 'weak %.4g vec/s'%d['value'], 'ms', round(d['ms_per_step'],4), '| strong ms', round(s.get('ms_per_step',0),4), 'eff', round(s.get('efficiency_vs_n1',0),3), 'ok', d.get('global_topk_matches_merge_of_shards'), s.get('global_topk_matches_merge_of_shards'))
except Exception as e: print('$wl $ex FAILED', e)"
  done
done
if [ "$n" != 1 ] && [ -z "${SKIP_CHECK:-}" ]; then
  cmd="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((port + 7)) tools/peer_check.py"
  if [ -n "$DRY" ]; then echo "$cmd"; else
    timeout 300 $cmd > gpurun_out/r2_peer_check_n$n.txt 2>&1
    grep -E "PEER_CHECK|us/step|MISMATCH" gpurun_out/r2_peer_check_n$n.txt | tail -14
  fi
fi
rm -f gpurun_out/tmp_line.json
