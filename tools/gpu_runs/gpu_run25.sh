#!/bin/bash
# ncu --set full capture of the headline kernel (FAST tier, 65 536 channels x 24 000 frames), after the plain run exited 0
set -u
O=gpurun_out; mkdir -p $O
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-small > $O/pre_ncu.json 2>/dev/null; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_bq_tma -s 3 -c 1 -f -o $O/k_bq_tma_fast_full python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-small > $O/ncu_bq.log 2>&1; echo "ncu rc=$?"
