#!/bin/bash
# C5 chunk-schedule sweep (needs a QZ_TUNING build of u8_batch.cu): candidate buffer size, growth margin, first chunk
for cfg in "4096 16 4" "4096 8 4" "4096 4 4" "4096 8 2" "8192 16 4" "8192 8 4" "8192 4 4" "8192 8 8" "16384 8 8"; do
  set -- $cfg
  echo -n "cap $1 margin $2 first cap/$3: "
  QZ_BATCH_CAP=$1 QZ_BATCH_MARGIN=$2 QZ_BATCH_FIRST_DIV=$3 REPS=8 timeout 60 python tools/batch_bench.py 2>&1 | tail -2 | tr "\n" " "; echo
done
