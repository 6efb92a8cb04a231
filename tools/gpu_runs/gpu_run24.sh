#!/bin/bash
# every secondary unit with the final kernels, one JSON line each (8 192 units x 48 000 frames unless stated)
set -u
O=gpurun_out; mkdir -p $O
: > $O/units_final.jsonl
for unit in filterfx delayfx fxrack dattorro chorus sampler; do
  timeout 400 python tools/bench_units.py $unit --cpu 2>/dev/null | grep '^{' >> $O/units_final.jsonl
done
for unit in filterfx delayfx fxrack sampler; do
  timeout 400 python tools/bench_units.py $unit --instances 65536 --frames 24000 2>/dev/null | grep '^{' >> $O/units_final.jsonl
done
timeout 400 python tools/bench_units.py voices --instances 65536 --frames 4800 --steps 10 2>/dev/null | grep '^{' >> $O/units_final.jsonl
python - <<'PY'
import json
for l in open('gpurun_out/units_final.jsonl'):
    d=json.loads(l); print(d['unit'], d.get('kernel'), d.get('instances', d.get('voices_per_gpu')), '%.3e'%d['value'], 'frac %.3f'%d.get('roofline',{}).get('frac',0), 'cpu', d.get('cpu_baseline',{}).get('value'))
PY
