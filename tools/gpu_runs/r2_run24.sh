#!/bin/bash
# round 2, run 24 (run 22 again: resting path with walked pointers and row reuse): chorus with the gen~ pitch shifter stage + stereo chorus (parity), timing
set -u
O=gpurun_out/r2_run24; mkdir -p $O
timeout 1200 python -m pytest tests/test_chorus_gpu.py tests/test_cpp_hosts_gpu.py tests/test_abi.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/pytest.log
timeout 900 python -m pytest tests/test_full_duration_gpu.py -m gpu -x -q -k chorus > $O/pytest_10s.log 2>&1; echo "10s rc=$?"; tail -3 $O/pytest_10s.log
for n in 4736 8192; do timeout 300 python tools/bench_units.py chorus --instances $n --frames 48000 --steps 3 2>/dev/null | grep '^{' | cut -c1-330; done | tee $O/chorus_units.jsonl
