#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_fx_gpu.py -m gpu -x -q > $O/pytest_fx.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_fx.log
: > $O/units2.jsonl
for u in filterfx delayfx fxrack; do
  timeout 300 python tools/bench_units.py $u >> $O/units2.jsonl
  for c in 32x4 64x4 64x8; do OLB_FX_TMA_CFG=$c timeout 300 python tools/bench_units.py $u >> $O/units2.jsonl; done
done
python - <<'PY'
import json
for l in open('gpurun_out/units2.jsonl'):
    d=json.loads(l); print(d['unit'], d.get('kernel'), d['env'], '%.3e'%d['value'], 'frac %.3f'%d['roofline']['frac'])
PY

timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_fx_tma -s 1 -c 1 -f -o $O/k_fx_tma_delay_full python tools/bench_units.py delayfx --frames 4800 --steps 1 > $O/ncu_fx.log 2>&1
ls -la $O | tail -8
