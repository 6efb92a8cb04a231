#!/bin/bash
# round 2, run 23: ncu --set full of k_pitch_taps (resting-phasor path), 8 192 instances x 8 192 frames
set -u
O=gpurun_out/r2_run23; mkdir -p $O
n=k_pitch_taps
timeout 300 python tools/bench_units.py chorus --instances 8192 --frames 8192 --steps 1 > $O/$n.plain.log 2>&1; echo "plain rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -f -k regex:k_pitch_taps -s 1 -c 1 -o $O/$n python tools/bench_units.py chorus --instances 8192 --frames 8192 --steps 1 > $O/$n.ncu.log 2>&1; echo "ncu rc=$?"
ncu -i $O/$n.ncu-rep --page details > $O/$n.details.txt 2>/dev/null
ncu -i $O/$n.ncu-rep --page raw --csv > $O/$n.raw.csv 2>/dev/null
ncu -i $O/$n.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip -9 > $O/$n.source_sass.csv.gz
