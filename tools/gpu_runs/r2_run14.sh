#!/bin/bash
# round 2, run 14: stereo fx parity; ncu source view of the rolled k_bq_skew tile at 4096 channels (lone warps)
set -u
O=gpurun_out/r2_run14; mkdir -p $O
timeout 900 python -m pytest tests/test_fx_gpu.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
for cfg in 48x32x4 192x8x4; do
  n=skew4096_$cfg
  OLB_BQ_SKEW_CFG=$cfg timeout 120 python tools/run_one.py 4096 48000 fast 4 1 > $O/$n.plain.log 2>&1; tail -1 $O/$n.plain.log
  OLB_BQ_SKEW_CFG=$cfg timeout 600 ncu --set full --clock-control none --import-source on -f -k regex:k_bq_skew -s 1 -c 1 -o $O/$n python tools/run_one.py 4096 48000 fast 4 1 > $O/$n.ncu.log 2>&1; echo "ncu rc=$?"
  ncu -i $O/$n.ncu-rep --page details > $O/$n.details.txt 2>/dev/null
  ncu -i $O/$n.ncu-rep --page raw --csv > $O/$n.raw.csv 2>/dev/null
  ncu -i $O/$n.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip -9 > $O/$n.source_sass.csv.gz
  rm -f $O/$n.ncu-rep
done
