#!/bin/bash
# round 2, run 37: setter race with the round-2 banks (plain and under ThreadSanitizer), odd-width stereo banks
set -u
O=gpurun_out/r2_run37; mkdir -p $O
timeout 900 python -m pytest tests/test_fx_gpu.py tests/test_cpp_hosts_gpu.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest.log
timeout 600 tests/cpp/setter_race 4096 400 64 > $O/setter_race.txt 2>&1; echo "race rc=$?"; cat $O/setter_race.txt
TSAN_OPTIONS="report_signal_unsafe=0 history_size=4" timeout 900 tests/cpp/setter_race_tsan 2048 150 64 > $O/setter_race_tsan.txt 2>&1; echo "tsan rc=$?"; grep -c "WARNING: ThreadSanitizer" $O/setter_race_tsan.txt; tail -3 $O/setter_race_tsan.txt
