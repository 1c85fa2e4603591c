#!/bin/bash
# Per workload: plain bench (must exit 0), then the ncu launch list and one --set full capture of the dominant kernel.
# Outputs under gpurun_out/prof_*; summaries are copied into profiles/ by hand.
set -u
for wl in u8 bin pq; do
  case $wl in u8|bin) kre=row_scan;; pq) kre=pq_scan2;; esac
  python bench.py --workload $wl --no-cpu --steps 20 --warmup 3 > gpurun_out/prof_${wl}_plain.json 2> gpurun_out/prof_${wl}_plain.err || { echo "$wl plain run failed"; continue; }
  ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/prof_${wl}_launches.csv \
      python bench.py --workload $wl --no-cpu --steps 5 --warmup 3 > /dev/null 2>&1
  ncu --set full --clock-control none --import-source on -k regex:$kre -s 3 -c 1 -f -o gpurun_out/prof_${wl}_full \
      python bench.py --workload $wl --no-cpu --steps 5 --warmup 3 > /dev/null 2>&1
  # keep text summaries only (gpurun_out/ is capped at 64 MiB)
  ncu -i gpurun_out/prof_${wl}_full.ncu-rep --page raw --csv > gpurun_out/prof_${wl}_raw.csv 2>/dev/null
  ncu -i gpurun_out/prof_${wl}_full.ncu-rep --page details > gpurun_out/prof_${wl}_details.txt 2>/dev/null
  rm -f gpurun_out/prof_${wl}_full.ncu-rep
  echo "$wl done"
done
