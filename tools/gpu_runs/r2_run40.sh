#!/bin/bash
# round 2, run 40: edge cases of every bank (empty blocks, one unit one frame at a time, rejected arguments)
set -u
O=gpurun_out/r2_run40; mkdir -p $O
timeout 900 python -m pytest tests/test_edges_gpu.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $O/pytest.log
