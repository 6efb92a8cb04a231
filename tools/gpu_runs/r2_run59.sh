#!/bin/bash
# round 2, run 59: k_dattorro_smem with the gathering taken out of the scan warp (4 warps: gather, scan, two frame warps): parity, then sizes
set -u
O=gpurun_out/r2_run59; mkdir -p $O
timeout 900 python -m pytest tests/test_dattorro_gpu.py tests/test_edges_gpu.py -m gpu -x -q > $O/pytest.log 2>&1; rc=$?; echo "pytest rc=$rc"; tail -3 $O/pytest.log
[ $rc -ne 0 ] && exit 1
timeout 900 python -m pytest tests/test_full_duration_gpu.py tests/test_fx_gpu.py -m gpu -x -q -k "dattorro or plate" > $O/pytest2.log 2>&1; echo "pytest2 rc=$?"; tail -3 $O/pytest2.log
for k in smem split; do for i in 148 296 1024 2560 8192; do
  OLB_PLATE_KERNEL=$k timeout 300 python tools/bench_units.py dattorro --instances $i --frames 48000 --steps 3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$k', d['instances'], '%.4e'%d['value'], '%.2f ms'%d['ms_per_step'], 'checksum', d['checksum'])"
done; done 2>&1 | tee $O/sizes.txt
