#!/bin/bash
# round 2, run 52: the continuous k_bq_roll with short tiles and deep rings (a warp holds two tiles around a boundary)
set -u
O=gpurun_out/r2_run52; mkdir -p $O
run() { OLB_BQ_SKEW_CFG=$2 timeout 120 python tools/run_one.py $1 $3 fast 4 20 2>&1 | tail -1 | sed "s/^/$2 /"; }
for cfg in 64x12x4c2 32x24x4c2 32x12x4c2 32x24x0c2 96x8x4c2; do run 8192 $cfg 24000; done 2>&1 | tee $O/sweep.txt
for cfg in 32x12x4c4 32x8x4c4 64x6x4c4 32x12x0c4; do run 16384 $cfg 24000; done 2>&1 | tee -a $O/sweep.txt
for F in 512 2400 4800; do run 8192 32x24x4c2 $F; done 2>&1 | tee -a $O/sweep.txt
