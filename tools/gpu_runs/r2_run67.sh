#!/bin/bash
# round 2, run 67: k_fx_split<RACK>, the reverb slot's mix moved from the wet warp (the heaviest) to the tail warp: parity, then both tiers
set -u
O=gpurun_out/r2_run67; mkdir -p $O
timeout 1200 python -m pytest tests/test_fx_gpu.py tests/test_full_duration_gpu.py tests/test_cpp_hosts_gpu.py -m gpu -x -q -k "not plate" > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
for i in 8192 65536; do for m in exact fma; do
  timeout 300 python tools/bench_units.py fxrack --instances $i --frames 48000 --steps 5 --mode $m >> $O/units.jsonl 2>> $O/units.err
done; done
python - <<'PY'
import json
for l in open('gpurun_out/r2_run67/units.jsonl'):
    d=json.loads(l); print(f"{d['unit']:9s} {d['mode']:5s} I={d['instances']:6d} {d['value']:.4e} inst-samp/s frac {d['roofline']['frac']:.3f} {d['kernel']} checksum {d['checksum']:.6f}")
PY
