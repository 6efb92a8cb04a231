#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_fx_gpu.py tests/test_full_duration_gpu.py tests/test_cpp_hosts_gpu.py -m gpu -x -q > $O/pytest_fx.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_fx.log
: > $O/units12.jsonl
for unit in delayfx fxrack; do
for inst in 8192 65536; do
  OLB_FX_SPLIT=0 timeout 300 python tools/bench_units.py $unit --instances $inst 2>/dev/null | grep '^{' | sed 's/^{/{"cfg": "one-warp", /' >> $O/units12.jsonl
  for cfg in 64x4 64x6 32x8 32x4; do
    OLB_FX_TMA_CFG=$cfg timeout 300 python tools/bench_units.py $unit --instances $inst 2>/dev/null | grep '^{' | sed "s/^{/{\"cfg\": \"split-$cfg\", /" >> $O/units12.jsonl
  done
done
done
python - <<'PY'
import json
for l in open('gpurun_out/units12.jsonl'):
    d=json.loads(l); print(d['cfg'], d['unit'], d.get('kernel'), d['instances'], '%.3e'%d['value'])
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_fx_split -s 1 -c 1 -f -o $O/k_rack_split_full python tools/bench_units.py fxrack --frames 4800 --steps 1 > $O/ncu_rack.log 2>&1
