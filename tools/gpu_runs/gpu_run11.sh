#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_wav_rack.py tests/test_fx_gpu.py tests/test_abi.py -x -q > $O/pytest_wav.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_wav.log
