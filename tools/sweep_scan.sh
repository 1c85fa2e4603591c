#!/bin/bash
# sweep of the scan ring geometry (experiments only)
for cfg in "24832 16 8" "24832 16 4" "49664 32 4" "49664 32 2"; do
  set -- $cfg
  echo "== tile_bytes=$1 align=$2 slots=$3"
  QZ_SCAN_TILE_BYTES=$1 QZ_SCAN_ROW_ALIGN=$2 QZ_SCAN_SLOTS=$3 python bench.py --steps 300 --no-cpu --workload ${WL:-u8} 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; print('step_ms', round(d['ms_per_step'],4), 'kernel_ms', round(r['kernel_ms'],4), 'GB/s', round(r['achieved'],1))"
done
