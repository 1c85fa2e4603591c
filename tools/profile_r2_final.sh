#!/bin/bash
# round 2, HEAD: bench lines of both arms, then (each after its plain command exited 0) the ncu launch list of the u8
# bench and one --set full capture of the u8 and of the binary scan kernel.  Text summaries only are kept.
set -u
python bench.py > gpurun_out/final_bench_u8.json 2> gpurun_out/final_bench_u8.err; echo "bench rc=$?"
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err; echo "reference arm rc=$?"
python bench.py --no-cpu --no-others --steps 5 --warmup 3 > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/final_launches_u8.csv \
    python bench.py --no-cpu --no-others --steps 5 --warmup 3 > /dev/null 2>&1
echo "launch list rc=$?"
for wl in u8 bin; do
  python bench.py --workload $wl --no-cpu --no-others --steps 5 --warmup 3 > /dev/null 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:row_scan -s 3 -c 1 -f -o gpurun_out/final_${wl}_full \
      python bench.py --workload $wl --no-cpu --no-others --steps 5 --warmup 3 > /dev/null 2>&1
  echo "$wl full rc=$?"
  ncu -i gpurun_out/final_${wl}_full.ncu-rep --page raw --csv > gpurun_out/final_${wl}_raw.csv 2>/dev/null
  ncu -i gpurun_out/final_${wl}_full.ncu-rep --page details > gpurun_out/final_${wl}_details.txt 2>/dev/null
  rm -f gpurun_out/final_${wl}_full.ncu-rep
done
