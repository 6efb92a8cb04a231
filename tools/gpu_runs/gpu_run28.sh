#!/bin/bash
# ncu --set full of the final fx pipeline kernel (FxRack, 8 192 instances, G = 2) after a plain run exited 0
set -u
O=gpurun_out; mkdir -p $O
python tools/bench_units.py fxrack --frames 4800 --steps 1 > /dev/null 2>&1; echo "plain rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_fx_split -s 1 -c 1 -f -o $O/k_fx_split_final python tools/bench_units.py fxrack --frames 4800 --steps 1 > $O/ncu_fx.log 2>&1; echo "ncu rc=$?"
