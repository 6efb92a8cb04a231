#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_voice_gpu.py tests/test_cpp_hosts_gpu.py -m gpu -x -q > $O/pytest_voice.log 2>&1; echo "pytest rc=$?"; tail -12 $O/pytest_voice.log
timeout 300 python tools/bench_units.py voices --instances 65536 --frames 4800 --steps 10 2>/dev/null | grep '^{' | tee $O/voices_v2.json | cut -c1-200
