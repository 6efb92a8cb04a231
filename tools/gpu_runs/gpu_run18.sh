#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_sampler_gpu.py -m gpu -x -q > $O/pytest_sampler.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_sampler.log
: > $O/units18.jsonl
timeout 300 python tools/bench_units.py sampler --cpu 2>/dev/null | grep '^{' >> $O/units18.jsonl
timeout 300 python tools/bench_units.py sampler --instances 65536 --frames 24000 2>/dev/null | grep '^{' >> $O/units18.jsonl
python - <<'PY'
import json
for l in open('gpurun_out/units18.jsonl'):
    d=json.loads(l); print(d['unit'], d.get('kernel'), d['instances'], '%.3e'%d['value'], 'frac %.3f'%d['roofline']['frac'], d.get('cpu_baseline',{}).get('value'))
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_sampler -s 1 -c 1 -f -o $O/k_sampler_full python tools/bench_units.py sampler --instances 65536 --frames 4800 --steps 1 > $O/ncu_sampler.log 2>&1
