#!/bin/bash
# round 2, run 55: ncu launch list of the bench command (gpu__time_duration.sum, one pass), after the same command exited 0 without ncu
set -u
O=gpurun_out/r2_run55; mkdir -p $O
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-seconds 0.3 --sustain-seconds 0.2"
timeout 600 $CMD > $O/bench_plain.json 2> $O/bench_plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv $CMD > $O/ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections, json
rows=[r for r in csv.reader(open('gpurun_out/r2_run55/launches.csv')) if len(r)>5]
hdr=rows[0]; ik=hdr.index('Kernel Name'); iv=hdr.index('Metric Value'); iu=hdr.index('Metric Unit')
t=collections.OrderedDict()
for r in rows[1:]:
    v=float(r[iv].replace(',','')); v*= {'ns':1e-3,'us':1,'ms':1e3}.get(r[iu],1)
    t.setdefault(r[ik][:70],[]).append(v)
tot=sum(sum(v) for v in t.values())
for k,v in t.items(): print(f"{k:72s} n={len(v):4d} mean {sum(v)/len(v):10.1f} us  total {sum(v)/1e3:9.3f} ms  {100*sum(v)/tot:5.1f} %")
d=json.load(open('gpurun_out/r2_run55/bench_plain.json')); print('event-timed ms_per_step', d['ms_per_step'], 'kernel', d['kernel'], 'gpu_launches', d['gpu_launches'])
PY
# ---- probe: is warp 0's gain product what sets the pace of k_bq_roll's four-warp pipeline?  (libolbank_nogain.so: wrong results, timing only)
cp ol_dsp_b200/libolbank.so /tmp/new.so
for lib in new nogain new nogain; do
  [ $lib = nogain ] && cp ol_dsp_b200/libolbank_nogain.so ol_dsp_b200/libolbank.so || cp /tmp/new.so ol_dsp_b200/libolbank.so
  for C in 8192 16384; do timeout 120 python tools/run_one.py $C 24000 fast 0 20 2>&1 | tail -1 | sed "s/^/$lib /"; done
done 2>&1 | tee $O/nogain_probe.txt
cp /tmp/new.so ol_dsp_b200/libolbank.so
# ---- ncu --set full: k_bq_roll with four sections per warp (16 384 channels), k_fx_tma<FILTER> in the fused tier
NCU="ncu --set full --clock-control none --import-source on -f"
cap() {  # cap <name> <kernel regex> <skip> -- command...
  n=$1; k=$2; s=$3; shift 3
  timeout 300 "$@" > $O/$n.plain.log 2>&1; rc=$?; echo "$n plain rc=$rc"; tail -1 $O/$n.plain.log | cut -c1-250
  [ $rc -eq 0 ] && timeout 600 $NCU -k regex:$k -s $s -c 1 -o $O/$n "$@" > $O/$n.ncu.log 2>&1; echo "$n ncu rc=$?"
  if [ -f $O/$n.ncu-rep ]; then
    ncu -i $O/$n.ncu-rep --page details > $O/$n.details.txt 2>/dev/null
    ncu -i $O/$n.ncu-rep --page raw --csv > $O/$n.raw.csv 2>/dev/null
    ncu -i $O/$n.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip -9 > $O/$n.source_sass.csv.gz
    rm -f $O/$n.ncu-rep
  fi
}
cap k_bq_roll_16384 k_bq_roll 1 python tools/run_one.py 16384 24000 fast 0 1
cap k_fx_filter_fma k_fx_tma 1 python tools/bench_units.py filterfx --instances 65536 --frames 12000 --steps 1 --mode fma
ls -la $O | head -30
