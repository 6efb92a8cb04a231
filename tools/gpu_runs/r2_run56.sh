#!/bin/bash
# round 2, run 56: k_fx_tma<FILTER> tile shapes at 65 536 and 8 192 instances, both tiers (residency: 64 x 4 is 32 KiB a CTA = 6 one-warp CTAs per SM)
set -u
O=gpurun_out/r2_run56; mkdir -p $O
for i in 65536 8192; do for m in exact fma; do for cfg in 64x4 32x4 32x8 64x8; do
  OLB_FX_TMA_CFG=$cfg timeout 300 python tools/bench_units.py filterfx --instances $i --frames 24000 --steps 5 --mode $m >> $O/units.jsonl 2>> $O/units.err
done; done; done
python - <<'PY'
import json
for l in open('gpurun_out/r2_run56/units.jsonl'):
    d=json.loads(l); print(f"{d['unit']:9s} {d['mode']:5s} I={d['instances']:6d} {d['env'].get('OLB_FX_TMA_CFG')} {d['value']:.4e} inst-samp/s frac {d['roofline']['frac']:.3f} {d['kernel']}")
PY
tail -3 $O/units.err
