#!/bin/bash
# round 2, run 32: k_bq_sys (two sections per warp, packed lanes) at 65 536 channels with shapes that keep all 1 024 CTAs resident
set -u
O=gpurun_out/r2_run32; mkdir -p $O
for mode in fast fma; do
  timeout 120 python tools/run_one.py 65536 24000 $mode 0 5 2>&1 | tail -1
  for cfg in 2x32x6x2 2x32x4x2 2x32x3x2 2x64x2x2 2x16x7x2; do OLB_BQ_SYS_CFG=$cfg timeout 120 python tools/run_one.py 65536 24000 $mode 3 5 2>&1 | tail -1 | sed "s/^/cfg=$cfg /"; done
done | tee $O/sys_65536.txt
