#!/bin/bash
# round 2, run 44: k_bq_roll variants (store lag 0/4, one or two blocks per trip) timed; ncu --set full of the lone-warp shape at 8 192 channels
set -u
O=gpurun_out/r2_run44; mkdir -p $O
run() { OLB_BQ_SKEW_CFG=$2 timeout 120 python tools/run_one.py $1 24000 fast 4 10 2>&1 | tail -1 | sed "s/^/$2 /"; }
for C in 8192 16384; do
  echo "--- $C channels"
  timeout 120 python tools/run_one.py $C 24000 fast 0 10 2>&1 | tail -1 | sed "s/^/auto /"
  for cfg in 128x6x0c2 128x6x1c2 128x6x4c2 128x6x5c2 96x8x4c2 96x8x5c2 64x6x0c2 64x6x4c2 64x6x5c2 64x6x0c4 64x6x4c4 64x6x5c4 128x3x4c4 128x3x5c4; do run $C $cfg; done
done 2>&1 | tee $O/sweep.txt
NCU="ncu --set full --clock-control none --import-source on -f"
cap() {  # cap <name> <kernel regex> <skip> -- command...
  n=$1; k=$2; s=$3; shift 3
  timeout 600 $NCU -k regex:$k -s $s -c 1 -o $O/$n "$@" > $O/$n.ncu.log 2>&1; echo "$n ncu rc=$?"
  if [ -f $O/$n.ncu-rep ]; then
    ncu -i $O/$n.ncu-rep --page details > $O/$n.details.txt 2>/dev/null
    ncu -i $O/$n.ncu-rep --page raw --csv > $O/$n.raw.csv 2>/dev/null
    ncu -i $O/$n.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip -9 > $O/$n.source_sass.csv.gz
    [ $(stat -c %s $O/$n.ncu-rep) -gt 8000000 ] && rm -f $O/$n.ncu-rep
  fi
}
OLB_BQ_SKEW_CFG=128x6x4c2 cap k_bq_roll_8192 k_bq_roll 1 python tools/run_one.py 8192 24000 fast 4 1
ls -la $O
