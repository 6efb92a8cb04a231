#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_voice -s 1 -c 1 -f -o $O/k_voice_full python tools/bench_units.py voices --instances 65536 --frames 2400 --steps 1 > $O/ncu_voice.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_dattorro_pf -s 1 -c 1 -f -o $O/k_dattorro_pf_full python tools/bench_units.py dattorro --frames 4800 --steps 1 > $O/ncu_dattorro.log 2>&1
tail -2 $O/ncu_voice.log $O/ncu_dattorro.log | cut -c1-300
