#!/bin/bash
# round 2, run 42: the fx banks' fused tier (olb_fx_set_mode(OLB_MODE_FMA)): parity of both tiers, then the C5 units in both
set -u
O=gpurun_out/r2_run42; mkdir -p $O
timeout 1500 python -m pytest tests/test_fx_gpu.py tests/test_edges_gpu.py tests/test_zz_no_alloc.py tests/test_full_duration_gpu.py -m gpu -x -q -s > $O/pytest_fx.log 2>&1; echo "fx pytest rc=$?"
grep -a "FMA - EXACT" $O/pytest_fx.log; tail -4 $O/pytest_fx.log
for u in filterfx delayfx fxrack; do for i in 8192 65536; do for m in exact fma; do
  timeout 300 python tools/bench_units.py $u --instances $i --frames 48000 --steps 5 --mode $m >> $O/units.jsonl 2>> $O/units.err
done; done; done
python - <<'PY'
import json
for l in open('gpurun_out/r2_run42/units.jsonl'):
    d=json.loads(l); print(f"{d['unit']:9s} {d['mode']:5s} I={d['instances']:6d} {d['value']:.4e} inst-samp/s frac {d['roofline']['frac']:.3f} {d['kernel']} checksum {d['checksum']:.6f}")
PY
tail -3 $O/units.err
