#!/bin/bash
# round 2, run 33: probe -- whole-cascade warps and fine-grained skew units concurrently on two streams
set -u
O=gpurun_out/r2_run33; mkdir -p $O
timeout 900 python tools/hybrid_probe.py 24000 2>&1 | tee $O/hybrid_probe.txt | tail -20
