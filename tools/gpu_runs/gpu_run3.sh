#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_chorus_gpu.py tests/test_fx_gpu.py tests/test_abi.py -m gpu -x -q > $O/pytest_chorus.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_chorus.log
: > $O/units3.jsonl
timeout 300 python tools/bench_units.py chorus --cpu >> $O/units3.jsonl
for n in 8192 32768 65536; do for u in filterfx delayfx fxrack; do timeout 300 python tools/bench_units.py $u --instances $n --frames 12000 >> $O/units3.jsonl; done; done
python - <<'PY'
import json
for l in open('gpurun_out/units3.jsonl'):
    d=json.loads(l); print(d['unit'], d.get('kernel'), d['instances'], '%.3e'%d['value'], 'frac %.3f'%d['roofline']['frac'], d.get('cpu_baseline',{}).get('value'))
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_chorus -s 1 -c 1 -f -o $O/k_chorus_full python tools/bench_units.py chorus --frames 4800 --steps 1 > $O/ncu_chorus.log 2>&1
