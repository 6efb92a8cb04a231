#!/bin/bash
# round 2, run 65: ncu --set full of k_dattorro_smem at HEAD (scan inputs fetched inside the chain), 148 instances = one plate per SM
set -u
O=gpurun_out/r2_run65; mkdir -p $O
NCU="ncu --set full --clock-control none --import-source on -f"
n=k_dattorro_smem
OLB_PLATE_KERNEL=smem timeout 600 $NCU -k regex:k_dattorro_smem -s 1 -c 1 -o $O/$n python tools/bench_units.py dattorro --instances 148 --frames 24000 --steps 1 > $O/$n.ncu.log 2>&1; echo "ncu rc=$?"
ncu -i $O/$n.ncu-rep --page details > $O/$n.details.txt 2>/dev/null
ncu -i $O/$n.ncu-rep --page raw --csv > $O/$n.raw.csv 2>/dev/null
ncu -i $O/$n.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip -9 > $O/$n.source_sass.csv.gz
rm -f $O/$n.ncu-rep; ls -la $O
