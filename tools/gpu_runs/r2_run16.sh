#!/bin/bash
# round 2, run 16: the no-allocation entry point (alone in a fresh process, then after the whole GPU suite), smoke
set -u
O=gpurun_out/r2_run16; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest_zz_alone.log 2>&1; echo "zz alone rc=$?"; tail -4 $O/pytest_zz_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -6 $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/smoke.log
