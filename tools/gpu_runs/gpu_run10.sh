#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_dattorro_gpu.py tests/test_cpp_hosts_gpu.py -m gpu -x -q > $O/pytest_plate.log 2>&1; echo "pytest rc=$?"; tail -12 $O/pytest_plate.log
: > $O/units10.jsonl
timeout 300 python tools/bench_units.py dattorro --cpu 2>/dev/null | grep '^{' >> $O/units10.jsonl
OLB_PLATE_KERNEL=pf timeout 300 python tools/bench_units.py dattorro 2>/dev/null | grep '^{' >> $O/units10.jsonl
timeout 300 python tools/bench_units.py dattorro --instances 16384 --frames 24000 2>/dev/null | grep '^{' >> $O/units10.jsonl
python - <<'PY'
import json
for l in open('gpurun_out/units10.jsonl'):
    d=json.loads(l); print(d['unit'], d['instances'], d.get('env'), '%.3e'%d['value'], 'frac %.3f'%d['roofline']['frac'], 'ring GB/s %.0f'%d['roofline']['ring_traffic_gbs'], d.get('cpu_baseline',{}).get('value'))
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_dattorro_rp -s 1 -c 1 -f -o $O/k_dattorro_rp_full python tools/bench_units.py dattorro --frames 4800 --steps 1 > $O/ncu_dattorro.log 2>&1
