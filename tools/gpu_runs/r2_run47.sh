#!/bin/bash
# round 2, run 47: k_bq_skew's rolled tile on the all-on rows of the stream's end tiles too: all biquad parity, then widths and block lengths
set -u
O=gpurun_out/r2_run47; mkdir -p $O
timeout 1500 python -m pytest tests/test_biquad_gpu.py tests/test_full_duration_gpu.py tests/test_edges_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
for C in 2048 4096 8192 12288 16384; do for F in 2400 4800 24000 48000; do timeout 120 python tools/run_one.py $C $F fast 0 20 2>&1 | tail -1; done; done 2>&1 | tee $O/auto.txt
for cfg in 64x6x4c4 64x6x5c4 64x6x0c4 128x3x5c4; do OLB_BQ_SKEW_CFG=$cfg timeout 120 python tools/run_one.py 16384 24000 fast 4 20 2>&1 | tail -1 | sed "s/^/$cfg /"; done 2>&1 | tee -a $O/auto.txt
