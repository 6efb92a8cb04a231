#!/bin/bash
# round 2, run 70: k_fx_split, the per-type switch hoisted out of the tile loop: all four cases at 8 192 instances, masks 0 and 3
set -u
O=gpurun_out/r2_run70; mkdir -p $O
for t in 0 3 1 2; do for u in delayfx fxrack; do for m in exact fma; do
  OLB_FX_SPLIT_TYPES=$t timeout 300 python tools/bench_units.py $u --instances 8192 --frames 48000 --steps 5 --mode $m >> $O/units.jsonl 2>> $O/units.err
done; done; done
python - <<'PY'
import json,collections
acc=collections.defaultdict(list)
for l in open('gpurun_out/r2_run70/units.jsonl'):
    d=json.loads(l); acc[(d['unit'],d['mode'],d['env'].get('OLB_FX_SPLIT_TYPES'))].append((d['roofline']['frac'], d['checksum']))
for k in sorted(acc): print(k, ' '.join('%.3f (%.6f)'%v for v in acc[k]))
PY
