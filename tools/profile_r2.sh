#!/bin/bash
# round 2: one --set full capture of the PQ scan and of the large-chunk batched kernel at HEAD (each after the
# same command exited 0 without ncu).  Reports are kept (read on the build machine with ncu -i).
set -u
python bench.py --workload pq --no-cpu --steps 5 --warmup 3 > gpurun_out/prof2_pq_plain.json 2> gpurun_out/prof2_pq_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:pq_scan2 -s 3 -c 1 -f -o gpurun_out/prof2_pq_full \
    python bench.py --workload pq --no-cpu --steps 5 --warmup 3 > /dev/null 2>&1
echo "pq rc=$?"
python bench.py --workload u8batch --no-cpu --steps 3 --warmup 3 > gpurun_out/prof2_u8batch_plain.json 2> gpurun_out/prof2_u8batch_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:u8_batch_kernel -s 13 -c 1 -f -o gpurun_out/prof2_u8batch_full \
    python bench.py --workload u8batch --no-cpu --steps 3 --warmup 3 > /dev/null 2>&1
echo "u8batch rc=$?"
ls -la gpurun_out/*.ncu-rep
