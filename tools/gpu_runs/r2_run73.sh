#!/bin/bash
# round 2, run 73: k_fx_split, the delay's wet warp with the type switch inside the tile loop, the rack's with it outside: parity, then 8 192 instances
set -u
O=gpurun_out/r2_run73; mkdir -p $O
timeout 1500 python -m pytest tests/test_fx_gpu.py tests/test_full_duration_gpu.py tests/test_cpp_hosts_gpu.py tests/test_edges_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
for rep in 1 2; do for u in delayfx fxrack; do for m in exact fma; do
  timeout 300 python tools/bench_units.py $u --instances 8192 --frames 48000 --steps 5 --mode $m >> $O/units.jsonl 2>> $O/units.err
done; done; done
python - <<'PY'
import json,collections
acc=collections.defaultdict(list)
for l in open('gpurun_out/r2_run73/units.jsonl'):
    d=json.loads(l); acc[(d['unit'],d['mode'])].append((d['roofline']['frac'], d['checksum']))
for k in sorted(acc): print(k, ' '.join('%.3f (%.6f)'%v for v in acc[k]))
PY
