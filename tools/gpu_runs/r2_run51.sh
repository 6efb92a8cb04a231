#!/bin/bash
# round 2, run 51: k_bq_roll as one continuous rolled run across tile boundaries: parity (under a short timeout), then timings
set -u
O=gpurun_out/r2_run51; mkdir -p $O
timeout 300 python -m pytest tests/test_biquad_gpu.py -m gpu -x -q -k "channel_pair" > $O/pytest_roll.log 2>&1; rc=$?; echo "roll pytest rc=$rc"; tail -5 $O/pytest_roll.log
[ $rc -ne 0 ] && exit 1
timeout 900 python -m pytest tests/test_full_duration_gpu.py -m gpu -x -q -k "wide_banks" > $O/pytest_wide.log 2>&1; echo "wide pytest rc=$?"; tail -3 $O/pytest_wide.log
run() { OLB_BQ_SKEW_CFG=$2 timeout 120 python tools/run_one.py $1 $3 fast 4 20 2>&1 | tail -1 | sed "s/^/$2 /"; }
for cfg in 128x6x4c2 128x6x0c2 96x8x4c2 64x6x4c2 32x6x4c2; do run 8192 $cfg 24000; done 2>&1 | tee $O/sweep.txt
for cfg in 64x6x4c4 64x6x0c4 128x4x4c4 32x6x4c4 64x6x4c2; do run 16384 $cfg 24000; done 2>&1 | tee -a $O/sweep.txt
for F in 512 2400 4800; do timeout 120 python tools/run_one.py 8192 $F fast 0 40 2>&1 | tail -1; done 2>&1 | tee -a $O/sweep.txt
