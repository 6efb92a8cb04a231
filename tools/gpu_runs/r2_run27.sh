#!/bin/bash
# round 2, run 27: check of the tree -- whole GPU suite, smoke, both bench arms
set -u
O=gpurun_out/r2_run27; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -6 $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -4 $O/smoke.log
python bench.py > $O/bench_1gpu.json 2> $O/bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_run27/bench_1gpu.json'))
print('value %.4e frac %.3f kernel %s parity %s' % (d['value'], d['roofline']['frac'], d['kernel'], d['parity_check']['bit_exact']))
print('tiers', {k: round(v['roofline_frac'],3) for k,v in d['tiers'].items()}, 'sustained %.3f' % d['sustained']['roofline_frac'])
print('configs1 %.3e %s frac %.3f' % (d['configs1']['value'], d['configs1']['kernel'], d['configs1']['roofline_frac']))
print('e2e %.3e pcie_frac %.3f' % (d['e2e']['value'], d['e2e']['pcie_frac']), 'voices %.3e' % d['voices']['value'], 'cpu %.3e' % d['cpu_baseline']['value'])
print('traffic', d['roofline']['traffic'], d['roofline']['traffic_commit'], 'clocks', d['clocks'])
r=json.load(open('gpurun_out/r2_run27/bench_ref.json')); print('ref %.3e cores %s' % (r['value'], r['cpu_baseline']['cores']))
PY
