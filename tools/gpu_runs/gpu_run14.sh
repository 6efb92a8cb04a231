#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_fx_gpu.py tests/test_full_duration_gpu.py tests/test_cpp_hosts_gpu.py -m gpu -x -q > $O/pytest_fx.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_fx.log
: > $O/units14.jsonl
for unit in delayfx fxrack; do
for inst in 4736 8192 16384 65536; do
  timeout 300 python tools/bench_units.py $unit --instances $inst 2>/dev/null | grep '^{' | sed 's/^{/{"cfg": "auto", /' >> $O/units14.jsonl
  for g in 1 2 4; do
    OLB_FX_SPLIT_G=$g timeout 300 python tools/bench_units.py $unit --instances $inst 2>/dev/null | grep '^{' | sed "s/^{/{\"cfg\": \"G=$g\", /" >> $O/units14.jsonl
  done
  OLB_FX_SPLIT_G=2 OLB_FX_TMA_CFG=32x4 timeout 300 python tools/bench_units.py $unit --instances $inst 2>/dev/null | grep '^{' | sed "s/^{/{\"cfg\": \"G=2 32x4\", /" >> $O/units14.jsonl
done
done
python - <<'PY'
import json
for l in open('gpurun_out/units14.jsonl'):
    d=json.loads(l); print(d['cfg'], d['unit'], d.get('kernel'), d['instances'], '%.3e'%d['value'])
PY
