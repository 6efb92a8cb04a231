#!/bin/bash
# round 2, run 34: ncu launch list of the bench command (gpu__time_duration.sum, one pass), after the same command exited 0 without ncu
set -u
O=gpurun_out/r2_run34; mkdir -p $O
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-seconds 0.3 --sustain-seconds 0.2"
timeout 600 $CMD > $O/bench_plain.json 2> $O/bench_plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv $CMD > $O/ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections, json
rows=[r for r in csv.reader(open('gpurun_out/r2_run34/launches.csv')) if len(r)>5]
hdr=rows[0]; ik=hdr.index('Kernel Name'); iv=hdr.index('Metric Value'); iu=hdr.index('Metric Unit')
t=collections.OrderedDict()
for r in rows[1:]:
    v=float(r[iv].replace(',','')); v*= {'ns':1e-3,'us':1,'ms':1e3}.get(r[iu],1)
    t.setdefault(r[ik][:70],[]).append(v)
tot=sum(sum(v) for v in t.values())
for k,v in t.items(): print(f"{k:72s} n={len(v):4d} mean {sum(v)/len(v):10.1f} us  total {sum(v)/1e3:9.3f} ms  {100*sum(v)/tot:5.1f} %")
d=json.load(open('gpurun_out/r2_run34/bench_plain.json')); print('event-timed ms_per_step', d['ms_per_step'], 'kernel', d['kernel'], 'gpu_launches', d['gpu_launches'])
PY
