#!/bin/bash
# round 2, run 31: C++ mirrors of the stereo banks and the stereo chorus; 128-frame skew tiles at 8 192 channels
set -u
O=gpurun_out/r2_run31; mkdir -p $O
timeout 1200 python -m pytest tests/test_cpp_hosts_gpu.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest.log
for cfg in 96x8x4r 128x6x4r; do OLB_BQ_SKEW_CFG=$cfg timeout 120 python tools/run_one.py 8192 96000 fast 4 5 2>&1 | tail -1 | sed "s/^/cfg=$cfg /"; done | tee $O/skew_8192.txt
