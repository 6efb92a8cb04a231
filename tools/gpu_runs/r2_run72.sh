#!/bin/bash
# round 2, run 72: the whole GPU suite, smoke and the bench line at the final HEAD
set -u
O=gpurun_out/r2_run72; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
python bench.py > $O/bench_1gpu.json 2> $O/bench.err; echo "bench rc=$?"; tail -2 $O/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_run72/bench_1gpu.json'))
print('value %.4e frac %.3f sustained %s' % (d['value'], d['roofline']['frac'], d.get('sustained',{}).get('roofline_frac')))
print('configs1 %.4e %.3f %s' % (d['configs1']['value'], d['configs1']['roofline_frac'], d['configs1']['kernel']))
print('tiers', {k:(round(v.get('roofline_frac'),3) if isinstance(v,dict) else v) for k,v in d.get('tiers',{}).items()})
print('e2e', d['e2e'].get('value'), d['e2e'].get('pcie_frac'), 'parity', d.get('parity_check',{}).get('bit_exact'), 'cpu', d.get('cpu_baseline',{}).get('value'))
PY
