#!/bin/bash
# round 2, run 57: every secondary unit in one run at HEAD (the round-2 refresh of profiles/r01_run32_units_final.jsonl), with the CPU figure beside it
set -u
O=gpurun_out/r2_run57; mkdir -p $O
u() { timeout 600 python tools/bench_units.py "$@" >> $O/units.jsonl 2>> $O/units.err; }
for m in exact fma; do for unit in filterfx delayfx fxrack; do
  u $unit --instances 8192 --frames 48000 --steps 3 --mode $m $([ $m = exact ] && echo --cpu)
  u $unit --instances 65536 --frames 24000 --steps 3 --mode $m
done; done
u dattorro --instances 8192 --frames 48000 --steps 3 --cpu
u dattorro --instances 296 --frames 48000 --steps 3
u dattorro --instances 1024 --frames 48000 --steps 3
u voices --instances 65536 --frames 2400 --steps 5
u chorus --instances 8192 --frames 48000 --steps 3 --cpu
u sampler --instances 8192 --frames 48000 --steps 3 --cpu
u sampler --instances 65536 --frames 24000 --steps 3
python - <<'PY'
import json
for l in open('gpurun_out/r2_run57/units.jsonl'):
    d=json.loads(l); r=d.get('roofline',{})
    print(f"{d['unit']:9s} {d.get('mode','-'):5s} I={d.get('instances', d.get('voices_per_gpu',0)):6d} F={d['frames']:6d} {d['value']:.4e} {d['metric']} frac {r.get('frac',0):.3f} {d.get('kernel','')} cpu {d.get('cpu_baseline',{}).get('value',0):.3e}")
PY
tail -3 $O/units.err
