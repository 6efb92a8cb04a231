#!/bin/bash
# round 2, run 35: voice bank with device-resident event lists
set -u
O=gpurun_out/r2_run35; mkdir -p $O
timeout 1200 python -m pytest tests/test_voice_gpu.py tests/test_cpp_hosts_gpu.py tests/test_comm_gpu.py tests/test_full_duration_gpu.py -m gpu -x -q -k "voice or Voice or poly or cpp or comm or event" > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
timeout 300 python tools/bench_units.py voices --instances 65536 --frames 4800 --steps 10 2>/dev/null | grep '^{' | cut -c1-260
