#!/bin/bash
# round 2, run 25: wide + long parity (banks at BASELINE width across the delay-line and counter wraps); 256-frame skew tiles (a TMA box is at most 256 rows)
set -u
O=gpurun_out/r2_run26; mkdir -p $O
timeout 1500 python -m pytest tests/test_full_duration_gpu.py -m gpu -x -q --durations=12 -k "baseline_width" > $O/pytest_wide.log 2>&1; echo "pytest rc=$?"; tail -22 $O/pytest_wide.log
for cfg in 192x8x4r 256x6x4r; do OLB_BQ_SKEW_CFG=$cfg timeout 120 python tools/run_one.py 4096 192000 fast 4 5 2>&1 | tail -1 | sed "s/^/cfg=$cfg /"; done | tee $O/skew_384.txt
