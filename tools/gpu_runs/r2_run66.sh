#!/bin/bash
# round 2, run 66: ncu --set full of k_fx_split<RACK> and <DELAY> at 8 192 instances (who waits for whom in the warp pipeline)
set -u
O=gpurun_out/r2_run66; mkdir -p $O
NCU="ncu --set full --clock-control none --import-source on -f"
for u in fxrack delayfx; do
  n=k_fx_split_$u
  timeout 600 $NCU -k regex:k_fx_split -s 1 -c 1 -o $O/$n python tools/bench_units.py $u --instances 8192 --frames 12000 --steps 1 > $O/$n.ncu.log 2>&1; echo "$u ncu rc=$?"
  ncu -i $O/$n.ncu-rep --page details > $O/$n.details.txt 2>/dev/null
  ncu -i $O/$n.ncu-rep --page raw --csv > $O/$n.raw.csv 2>/dev/null
  ncu -i $O/$n.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip -9 > $O/$n.source_sass.csv.gz
  rm -f $O/$n.ncu-rep
done
ls -la $O
