#!/bin/bash
# round 2, run 43: k_bq_roll (two channels per lane, SPW sections per warp, rolled tile) and k_bq_quad: parity, then shapes at 8 192 / 16 384 / 12 288 channels
set -u
O=gpurun_out/r2_run43; mkdir -p $O
timeout 900 python -m pytest tests/test_biquad_gpu.py -m gpu -x -q -k "channel_pair or rolled_tile" > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
run() { OLB_BQ_SKEW_CFG=$2 timeout 120 python tools/run_one.py $1 24000 fast 4 10 2>&1 | tail -1 | sed "s/^/$2 /"; }
for C in 8192 16384 12288 6144; do
  echo "--- $C channels"
  timeout 120 python tools/run_one.py $C 24000 fast 0 10 2>&1 | tail -1 | sed "s/^/auto /"
  for cfg in 128x6x4c2 256x3x4c2 96x8x4c2 128x3x4c2 64x6x4c2 64x3x4c2 128x6x4c4 128x3x4c4 64x6x4c4 64x3x4c4 128x6x4q 128x4x4q 96x8x4q 48x8x4q; do run $C $cfg; done
done 2>&1 | tee $O/sweep.txt
