#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_chorus_gpu.py -m gpu -x -q > $O/pytest_chorus.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_chorus.log
timeout 300 python tools/bench_units.py chorus --cpu 2>/dev/null | grep '^{' > $O/units16.jsonl
timeout 300 python tools/bench_units.py chorus --instances 4736 2>/dev/null | grep '^{' >> $O/units16.jsonl
python - <<'PY'
import json
for l in open('gpurun_out/units16.jsonl'):
    d=json.loads(l); print(d['unit'], d.get('kernel'), d['instances'], '%.3e'%d['value'], 'frac %.3f'%d['roofline']['frac'], d.get('cpu_baseline',{}).get('value'))
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_chorus -s 1 -c 1 -f -o $O/k_chorus3_full python tools/bench_units.py chorus --frames 4800 --steps 1 > $O/ncu_chorus.log 2>&1
