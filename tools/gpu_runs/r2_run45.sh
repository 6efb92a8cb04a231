#!/bin/bash
# round 2, run 45: k_bq_roll in the automatic choice: biquad + full-duration + bench-contract tests, then the auto choice at several widths
set -u
O=gpurun_out/r2_run45; mkdir -p $O
timeout 1500 python -m pytest tests/test_biquad_gpu.py tests/test_full_duration_gpu.py tests/test_bench_contract.py tests/test_edges_gpu.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
for C in 4096 6144 8192 9472 12288 14336 16384 18944 24576; do
  timeout 120 python tools/run_one.py $C 24000 fast 0 10 2>&1 | tail -1
done 2>&1 | tee $O/auto.txt
for C in 8192 16384; do for F in 512 1024 2400 4800; do timeout 120 python tools/run_one.py $C $F fast 0 20 2>&1 | tail -1; done; done 2>&1 | tee -a $O/auto.txt
