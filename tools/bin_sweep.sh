#!/bin/bash
# binary scan: mapping / ring geometry / warm-up sweep (scores-only vs top-k kernel time)
run() { env WL=bin "$@" python tools/scan_ab.py 2>&1 | tail -1 | sed "s/^/$* : /"; }
run A=0
run QZ_SCAN_LPR=8
run QZ_SCAN_TILE_BYTES=9300
run QZ_SCAN_SLOTS=8 QZ_SCAN_GROUPS=4 QZ_SCAN_TILE_BYTES=24832
run QZ_SCAN_SLOTS=8 QZ_SCAN_GROUPS=8 QZ_SCAN_TILE_BYTES=24832
run QZ_SCAN_GROUPS=4
run QZ_SCAN_GROUPS=16
run QZ_SCAN_WARM=1024
