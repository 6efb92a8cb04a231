#!/bin/bash
# round 2, run 39: CC-map sweeps (every controller number) of the fx banks and the voice bank against the oracle
set -u
O=gpurun_out/r2_run39; mkdir -p $O
timeout 1500 python -m pytest tests/test_fx_gpu.py tests/test_voice_gpu.py -m gpu -x -q -k "sweep" --durations=6 > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -14 $O/pytest.log
