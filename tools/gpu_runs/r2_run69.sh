#!/bin/bash
# round 2, run 69: which warps of k_fx_split should pick the filter tap at compile time (OLB_FX_SPLIT_TYPES bit 0 wet, bit 1 tail)
set -u
O=gpurun_out/r2_run69; mkdir -p $O
for rep in 1 2; do for t in 0 1 2 3; do for u in delayfx fxrack; do for m in exact fma; do
  OLB_FX_SPLIT_TYPES=$t timeout 300 python tools/bench_units.py $u --instances 8192 --frames 48000 --steps 5 --mode $m >> $O/units.jsonl 2>> $O/units.err
done; done; done; done
python - <<'PY'
import json,collections
acc=collections.defaultdict(list)
for l in open('gpurun_out/r2_run69/units.jsonl'):
    d=json.loads(l); acc[(d['unit'],d['mode'],d['env'].get('OLB_FX_SPLIT_TYPES'))].append(d['roofline']['frac'])
for k in sorted(acc): print(k, ' '.join('%.3f'%v for v in acc[k]))
PY
