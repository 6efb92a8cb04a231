#!/bin/bash
# round 2, run 63: k_dattorro_smem, warp roles rotated in odd CTAs (two plates share an SM: where do the two chains sit?)
set -u
O=gpurun_out/r2_run63; mkdir -p $O
for rot in 0 1 2 3 4; do for i in 296 1024 8192; do
  OLB_PLATE_ROT=$rot OLB_PLATE_KERNEL=smem timeout 300 python tools/bench_units.py dattorro --instances $i --frames 48000 --steps 3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('rot $rot', d['instances'], '%.4e'%d['value'], '%.2f ms'%d['ms_per_step'], 'checksum', d['checksum'])"
done; done 2>&1 | tee $O/rot.txt
