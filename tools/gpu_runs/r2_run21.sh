#!/bin/bash
# round 2, run 21: chorus launch list after the taps kernel was batched (cut = 8 192 frames at 8 192 instances)
set -u
O=gpurun_out/r2_run21; mkdir -p $O
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/chorus_launches.csv python tools/bench_units.py chorus --instances 8192 --frames 16384 --steps 1 > $O/ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open('gpurun_out/r2_run21/chorus_launches.csv')) if len(r)>5]
hdr=rows[0]; ik=hdr.index('Kernel Name'); iv=hdr.index('Metric Value'); iu=hdr.index('Metric Unit')
t=collections.defaultdict(list)
for r in rows[1:]:
    v=float(r[iv].replace(',','')); v*= {'ns':1e-3,'us':1,'ms':1e3}.get(r[iu],1)
    t[r[ik][:40]].append(v)
for k,v in t.items(): print(f"{k:42s} n={len(v):4d} mean {sum(v)/len(v):9.1f} us  total {sum(v)/1e3:8.3f} ms")
PY
