#!/bin/bash
# consolidated check of the tree: full GPU suite, smoke, bench (both arms), launch list of the bench command
set -u
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_full.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_gpu_full.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/smoke.log
timeout 900 python bench.py > $O/bench_final.json 2> $O/bench_final.err; echo "bench rc=$?"; cat $O/bench_final.json
timeout 900 python bench.py --impl reference > $O/bench_ref_final.json 2> $O/bench_ref_final.err; echo "ref rc=$?"; cat $O/bench_ref_final.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_final.csv python bench.py --steps 2 --warmup 3 > $O/ncu_launches.log 2>&1; echo "ncu rc=$?"
