#!/bin/bash
# round-1 re-entry run: parity suite, headline bench, lane sweeps of the small banks, ncu evidence
set -u
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
python bench.py > $O/bench_1gpu.json 2> $O/bench_1gpu.err; echo "bench rc=$?"; cat $O/bench_1gpu.json
python bench.py --channels 4096 --block-frames 192000 --steps 30 --no-e2e --no-cpu-baseline > $O/bench_c2.json 2>> $O/bench_1gpu.err; cat $O/bench_c2.json
: > $O/units.jsonl
for u in filterfx delayfx fxrack; do
  OLB_FX_KERNEL=simple python tools/bench_units.py $u | sed 's/^/{"variant":"simple"} /' >> $O/units.jsonl
  for L in 32 16 8; do OLB_FX_LANES=$L python tools/bench_units.py $u | sed "s/^/{\"variant\":\"pf L=$L\"} /" >> $O/units.jsonl; done
done
OLB_PLATE_KERNEL=simple python tools/bench_units.py dattorro | sed 's/^/{"variant":"simple"} /' >> $O/units.jsonl
for L in 32 16 8; do OLB_PLATE_LANES=$L python tools/bench_units.py dattorro | sed "s/^/{\"variant\":\"pf L=$L\"} /" >> $O/units.jsonl; done
python tools/bench_units.py voices --instances 65536 --frames 4800 --steps 5 >> $O/units.jsonl
cut -c1-330 $O/units.jsonl
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 > $O/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_bq_tma -s 2 -c 1 -f -o $O/k_bq_tma_full python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > $O/ncu_full.log 2>&1
ls -la $O
