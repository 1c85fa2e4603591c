#!/bin/bash
# round 2: one --set full capture of a workload's dominant kernel at HEAD (after the same command exited 0 without
# ncu).  usage: tools/profile_r2.sh pq|u8batch ...   Reports are kept (read on the build machine with ncu -i).
set -u
for wl in "$@"; do
  case $wl in pq) kre=pq_scan2; skip=3; steps=5;; u8batch) kre=u8_batch_kernel; skip=13; steps=3;; u8|bin) kre=row_scan; skip=3; steps=5;; esac
  python bench.py --workload $wl --no-cpu --steps $steps --warmup 3 > gpurun_out/prof2_${wl}_plain.json 2> gpurun_out/prof2_${wl}_plain.err && \
  ncu --set full --clock-control none --import-source on -k regex:$kre -s $skip -c 1 -f -o gpurun_out/prof2_${wl}_full \
      python bench.py --workload $wl --no-cpu --steps $steps --warmup 3 > /dev/null 2>&1
  echo "$wl rc=$?"
done
ls -la gpurun_out/*.ncu-rep
