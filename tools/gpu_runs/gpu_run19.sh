#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_dattorro_gpu.py tests/test_full_duration_gpu.py tests/test_fx_gpu.py -m gpu -x -q -k "dattorro or plate or Plate" > $O/pytest_plate.log 2>&1; echo "pytest rc=$?"; tail -12 $O/pytest_plate.log
: > $O/units19.jsonl
for inst in 4736 8192 16384; do
  timeout 300 python tools/bench_units.py dattorro --instances $inst --frames 24000 2>/dev/null | grep '^{' | sed 's/^{/{"cfg": "split", /' >> $O/units19.jsonl
  OLB_PLATE_KERNEL=rp timeout 300 python tools/bench_units.py dattorro --instances $inst --frames 24000 2>/dev/null | grep '^{' | sed 's/^{/{"cfg": "rp", /' >> $O/units19.jsonl
done
python - <<'PY'
import json
for l in open('gpurun_out/units19.jsonl'):
    d=json.loads(l); print(d['cfg'], d['unit'], d['instances'], '%.3e'%d['value'])
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_dattorro_split -s 1 -c 1 -f -o $O/k_dattorro_split_full python tools/bench_units.py dattorro --frames 4800 --steps 1 > $O/ncu_plate.log 2>&1
