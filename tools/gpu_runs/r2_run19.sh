#!/bin/bash
# round 2, run 19: where the chorus time goes with the pitch stage (launch list), cut-length sweep
set -u
O=gpurun_out/r2_run19; mkdir -p $O
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/chorus_launches.csv python tools/bench_units.py chorus --instances 8192 --frames 12288 --steps 1 > $O/ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open('gpurun_out/r2_run19/chorus_launches.csv')) if len(r)>5]
hdr=rows[0]; ik=hdr.index('Kernel Name'); iv=hdr.index('Metric Value'); iu=hdr.index('Metric Unit')
t=collections.defaultdict(list)
for r in rows[1:]:
    v=float(r[iv].replace(',','')); v*= {'ns':1e-3,'us':1,'ms':1e3}.get(r[iu],1)
    t[r[ik][:40]].append(v)
for k,v in t.items(): print(f"{k:42s} n={len(v):4d} mean {sum(v)/len(v):9.1f} us  total {sum(v)/1e3:8.3f} ms")
PY
for cut in 512 1536 4096 12288 48000; do OLB_CHORUS_CUT=$cut timeout 300 python tools/bench_units.py chorus --instances 8192 --frames 48000 --steps 3 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cut $cut', '%.3e'%d['value'], '%.3f ms'%d['ms_per_step'])"; done | tee $O/cut_sweep.txt
