#!/bin/bash
# round 2, run 68: k_fx_split with the filter type picked at compile time when a warp's instances share it (wet and tail warps): parity, then both ring kinds in both tiers
set -u
O=gpurun_out/r2_run68; mkdir -p $O
timeout 1500 python -m pytest tests/test_fx_gpu.py tests/test_full_duration_gpu.py tests/test_cpp_hosts_gpu.py tests/test_edges_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
for u in delayfx fxrack; do for i in 8192 65536; do for m in exact fma; do
  timeout 300 python tools/bench_units.py $u --instances $i --frames 48000 --steps 5 --mode $m >> $O/units.jsonl 2>> $O/units.err
done; done; done
python - <<'PY'
import json
for l in open('gpurun_out/r2_run68/units.jsonl'):
    d=json.loads(l); print(f"{d['unit']:9s} {d['mode']:5s} I={d['instances']:6d} {d['value']:.4e} inst-samp/s frac {d['roofline']['frac']:.3f} {d['kernel']} checksum {d['checksum']:.6f}")
PY
