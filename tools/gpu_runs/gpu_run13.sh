#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
: > $O/units13.jsonl
for unit in delayfx fxrack filterfx; do
for inst in 4736 9472; do
  OLB_FX_SPLIT=0 timeout 300 python tools/bench_units.py $unit --instances $inst 2>/dev/null | grep '^{' | sed 's/^{/{"cfg": "one-warp", /' >> $O/units13.jsonl
  for cfg in 64x4 32x4; do
    OLB_FX_TMA_CFG=$cfg timeout 300 python tools/bench_units.py $unit --instances $inst 2>/dev/null | grep '^{' | sed "s/^{/{\"cfg\": \"split-$cfg\", /" >> $O/units13.jsonl
  done
done
done
python - <<'PY'
import json
for l in open('gpurun_out/units13.jsonl'):
    d=json.loads(l); print(d['cfg'], d['unit'], d.get('kernel'), d['instances'], '%.3e'%d['value'], 'cycles/frame @1.9GHz %.0f'%(1.9e9*d['instances']/d['value']))
PY
