#!/bin/bash
# round 2, run 71: parity of the fx banks with the final k_fx_split (type switch hoisted, wet warp specialised), every mask value
set -u
O=gpurun_out/r2_run71; mkdir -p $O
timeout 1500 python -m pytest tests/test_fx_gpu.py tests/test_full_duration_gpu.py tests/test_cpp_hosts_gpu.py tests/test_edges_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
OLB_FX_SPLIT_TYPES=3 timeout 900 python -m pytest tests/test_fx_gpu.py -m gpu -x -q -k "variant or large_bank or golden or fma_tier_every" > $O/pytest_mask3.log 2>&1; echo "mask3 rc=$?"; tail -2 $O/pytest_mask3.log
OLB_FX_SPLIT_TYPES=0 timeout 900 python -m pytest tests/test_fx_gpu.py -m gpu -x -q -k "variant or large_bank" > $O/pytest_mask0.log 2>&1; echo "mask0 rc=$?"; tail -2 $O/pytest_mask0.log
