#!/bin/bash
# round 2, run 58: the C++ host tests with the fx banks' mode setter in the control thread's mix; the same race under ThreadSanitizer
set -u
O=gpurun_out/r2_run58; mkdir -p $O
timeout 900 python -m pytest tests/test_cpp_hosts_gpu.py -m gpu -x -q > $O/pytest_cpp.log 2>&1; echo "cpp pytest rc=$?"; tail -3 $O/pytest_cpp.log
TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0" timeout 600 tests/cpp/setter_race_tsan 4096 200 64 > $O/tsan.out 2> $O/tsan.err; echo "tsan rc=$?"
tail -2 $O/tsan.out; grep -c "WARNING: ThreadSanitizer" $O/tsan.err; grep "WARNING: ThreadSanitizer" -A12 $O/tsan.err | head -40
