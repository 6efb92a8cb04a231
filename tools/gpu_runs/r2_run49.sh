#!/bin/bash
# round 2, run 49: A/B on ONE box -- the library of commit 601fdfb (libolbank_head.so) against the working tree -- for the banks whose
# kernel did not change on purpose (12 288 / 24 576 channels: k_bq_skew's unrolled tile; 65 536: k_bq_tma), then the bench line
set -u
O=gpurun_out/r2_run49; mkdir -p $O
cp ol_dsp_b200/libolbank.so /tmp/new.so; cp ol_dsp_b200/libolbank_head.so /tmp/head.so
for rep in 1 2; do for lib in head new; do
  cp /tmp/$lib.so ol_dsp_b200/libolbank.so
  for C in 12288 24576 4096 65536; do timeout 120 python tools/run_one.py $C 24000 fast 0 20 2>&1 | tail -1 | sed "s/^/$lib /"; done
done; done 2>&1 | tee $O/ab.txt
cp /tmp/new.so ol_dsp_b200/libolbank.so
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,temperature.gpu --format=csv | tee -a $O/ab.txt
python bench.py --steps 20 > $O/bench_1gpu.json 2> $O/bench.err; echo "bench rc=$?"; tail -2 $O/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_run49/bench_1gpu.json'))
print('value %.4e frac %.3f sustained %s' % (d['value'], d['roofline']['frac'], d.get('sustained',{}).get('roofline_frac')))
print('configs1', d['configs1']); print('strong', d.get('strong')); print('tiers', {k:(v.get('roofline_frac') if isinstance(v,dict) else v) for k,v in d.get('tiers',{}).items()})
print('e2e', d['e2e'].get('value'), d['e2e'].get('pcie_frac'), 'parity', d.get('parity_check'))
PY
