#!/bin/bash
# round 2, run 15: rolled-tile routing (parity at every tile height, 10 s wide banks), bench line
set -u
O=gpurun_out/r2_run15; mkdir -p $O
timeout 1500 python -m pytest tests/test_biquad_gpu.py tests/test_full_duration_gpu.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
python bench.py --steps 50 > $O/bench_1gpu.json 2> $O/bench.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_run15/bench_1gpu.json'))
print(d['value'], d['roofline']['frac'], d['configs1'], d['strong'], d['parity_check']['bit_exact'])
PY
for c in "4096 192000" "2048 192000" "1024 192000" "8192 96000" "12288 48000" "16384 48000"; do set -- $c; timeout 120 python tools/run_one.py $1 $2 fast 0 5 2>&1 | tail -1; done | tee $O/auto_choice.txt
