#!/bin/bash
# round 2, run 50: after restoring the unrolled tiles' surroundings: all biquad parity; A/B against commit 601fdfb's library on one box
set -u
O=gpurun_out/r2_run50; mkdir -p $O
timeout 1500 python -m pytest tests/test_biquad_gpu.py tests/test_full_duration_gpu.py tests/test_edges_gpu.py tests/test_bench_contract.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
cp ol_dsp_b200/libolbank.so /tmp/new.so; cp ol_dsp_b200/libolbank_head.so /tmp/head.so
for lib in head new head new; do
  cp /tmp/$lib.so ol_dsp_b200/libolbank.so
  for C in 2048 4096 8192 12288 16384 24576; do timeout 120 python tools/run_one.py $C 24000 fast 0 20 2>&1 | tail -1 | sed "s/^/$lib /"; done
  for C in 4096 8192; do timeout 120 python tools/run_one.py $C 2400 fast 0 40 2>&1 | tail -1 | sed "s/^/$lib /"; done
done 2>&1 | tee $O/ab.txt
cp /tmp/new.so ol_dsp_b200/libolbank.so
