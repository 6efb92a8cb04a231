#!/bin/bash
# round 2, run 12: rolled FAST tile of k_bq_skew (parity + timing), stereo fx banks (parity)
set -u
O=gpurun_out/r2_run13; mkdir -p $O
timeout 900 python -m pytest tests/test_biquad_gpu.py tests/test_fx_gpu.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
for c in "4096 192000" "8192 96000" "16384 48000" "32768 48000"; do
  set -- $c
  for v in 0 3 4; do timeout 120 python tools/run_one.py $1 $2 fast $v 5 2>&1 | tail -1; done
  for cfg in 48x32x4 96x16x4 192x8x4 192x4x4 96x8x4 48x16x4 24x8x4 48x8x4; do OLB_BQ_SKEW_CFG=$cfg timeout 120 python tools/run_one.py $1 $2 fast 4 5 2>&1 | tail -1 | sed "s/^/cfg=$cfg /"; done
done | tee $O/skew_sweep.txt
