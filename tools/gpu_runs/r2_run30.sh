#!/bin/bash
# round 2, run 30: unrolled skew tile with its original gain handling (16 384 / 32 768 channels), dense event list test
set -u
O=gpurun_out/r2_run30; mkdir -p $O
timeout 1200 python -m pytest tests/test_biquad_gpu.py tests/test_voice_gpu.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest.log
for c in "16384 96000" "16384 48000" "12288 48000" "24576 48000" "8192 96000" "4096 192000"; do set -- $c; timeout 120 python tools/run_one.py $1 $2 fast 0 5 2>&1 | tail -1; done | tee $O/auto_choice.txt
