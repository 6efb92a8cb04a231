#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_fx_gpu.py tests/test_abi.py -m gpu -x -q > $O/pytest_fx.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_fx.log
: > $O/units5.jsonl
for n in 8192 65536; do for u in delayfx fxrack; do timeout 300 python tools/bench_units.py $u --instances $n --frames 12000 2>/dev/null | grep '^{' >> $O/units5.jsonl; done; done

python - <<'PY'
import json
for l in open('gpurun_out/units5.jsonl'):
    d=json.loads(l); print(d['unit'], d.get('kernel'), d['instances'], d.get('env'), '%.3e'%d['value'], 'frac %.3f'%d['roofline']['frac'])
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_fx_tma -s 1 -c 1 -f -o $O/k_fx_tma_delay_v3 python tools/bench_units.py delayfx --frames 4800 --steps 1 > $O/ncu_fx.log 2>&1
