#!/bin/bash
# round 2, run 46: k_bq_roll with the fast path on the all-on rows of the stream's end tiles: parity, short blocks
set -u
O=gpurun_out/r2_run46; mkdir -p $O
timeout 1500 python -m pytest tests/test_biquad_gpu.py tests/test_full_duration_gpu.py -m gpu -x -q -k "channel_pair or wide_banks or rolled or skewed" > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
for C in 8192 16384; do for F in 512 1024 2400 4800 24000; do timeout 120 python tools/run_one.py $C $F fast 0 20 2>&1 | tail -1; done; done 2>&1 | tee $O/auto.txt
for F in 512 1024 2400 4800; do OLB_BQ_SKEW_CFG=64x6x4c2 timeout 120 python tools/run_one.py 8192 $F fast 4 20 2>&1 | tail -1 | sed "s/^/64x6x4c2 /"; done 2>&1 | tee -a $O/auto.txt
for F in 512 1024 2400 4800; do OLB_BQ_SKEW_CFG=96x8x4r timeout 120 python tools/run_one.py 8192 $F fast 4 20 2>&1 | tail -1 | sed "s/^/skew 96x8x4r /"; done 2>&1 | tee -a $O/auto.txt
for F in 512 1024 2400 4800; do OLB_BQ_SKEW_CFG=48x16x4r timeout 120 python tools/run_one.py 8192 $F fast 4 20 2>&1 | tail -1 | sed "s/^/skew 48x16x4r /"; done 2>&1 | tee -a $O/auto.txt
