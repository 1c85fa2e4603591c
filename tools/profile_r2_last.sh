#!/bin/bash
# last verification of round 2 at HEAD (1 GPU): GPU suite, smoke, default bench line, then (after the plain command exited 0)
# the ncu launch list of the u8 bench and one --set full capture of the u8 scan kernel.  Text summaries only are kept.
set -u
timeout 200 python -m pytest tests -m gpu -x -q > gpurun_out/last_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/last_pytest_gpu.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/last_smoke.log 2>&1; echo "smoke rc=$?"
timeout 300 python bench.py > gpurun_out/last_bench_u8.json 2> gpurun_out/last_bench_u8.err; echo "bench rc=$?"
timeout 60 python bench.py --no-cpu --no-others --steps 5 --warmup 3 > /dev/null 2>&1 && \
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/last_launches_u8.csv \
    python bench.py --no-cpu --no-others --steps 5 --warmup 3 > /dev/null 2>&1
echo "launch list rc=$?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:row_scan -s 3 -c 1 -f -o gpurun_out/last_u8_full \
    python bench.py --no-cpu --no-others --steps 5 --warmup 3 > /dev/null 2>&1
echo "u8 full rc=$?"
ncu -i gpurun_out/last_u8_full.ncu-rep --page raw --csv > gpurun_out/last_u8_raw.csv 2>/dev/null
ncu -i gpurun_out/last_u8_full.ncu-rep --page details > gpurun_out/last_u8_details.txt 2>/dev/null
rm -f gpurun_out/last_u8_full.ncu-rep
