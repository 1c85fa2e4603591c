#!/bin/bash
# 1/2/4/8-GPU runs of bench.py exactly as the driver launches them (u8 headline workload, default exchange = peer memory)
for n in 1 2 4 8; do
  if [ $n = 1 ]; then timeout 200 python bench.py --gpus 1 --steps 300 --warmup 20 --no-cpu > gpurun_out/scale_final_n$n.log 2>&1
  else timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700+n)) bench.py --gpus $n --steps 300 --warmup 20 --no-cpu > gpurun_out/scale_final_n$n.log 2>&1; fi
  tail -1 gpurun_out/scale_final_n$n.log >> gpurun_out/scale_final.jsonl
  tail -1 gpurun_out/scale_final_n$n.log | cut -c1-200
done
