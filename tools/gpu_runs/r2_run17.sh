#!/bin/bash
# round 2, run 17 (re-run of run 11 after the machine that held its output was replaced): evidence at HEAD -- ncu --set full of the kernels VERDICT item 6 names
set -u
O=gpurun_out/r2_run17; mkdir -p $O
NCU="ncu --set full --clock-control none --import-source on -f"
cap() {  # cap <name> <kernel regex> <skip> -- command...
  n=$1; k=$2; s=$3; shift 3
  timeout 300 "$@" > $O/$n.plain.log 2>&1; rc=$?; echo "$n plain rc=$rc"; tail -1 $O/$n.plain.log | cut -c1-250
  [ $rc -eq 0 ] && timeout 600 $NCU -k regex:$k -s $s -c 1 -o $O/$n "$@" > $O/$n.ncu.log 2>&1; echo "$n ncu rc=$?"
  # the reports of the big unrolled kernels are ~30 MB each: keep text exports, drop reports over 8 MB
  if [ -f $O/$n.ncu-rep ]; then
    ncu -i $O/$n.ncu-rep --page details > $O/$n.details.txt 2>/dev/null
    ncu -i $O/$n.ncu-rep --page raw --csv > $O/$n.raw.csv 2>/dev/null
    ncu -i $O/$n.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip -9 > $O/$n.source_sass.csv.gz
    [ $(stat -c %s $O/$n.ncu-rep) -gt 8000000 ] && rm -f $O/$n.ncu-rep
  fi
}
cap k_voice        k_voice    1 python tools/bench_units.py voices   --instances 65536 --frames 2400  --steps 1
cap k_fx_filter    k_fx_tma   1 python tools/bench_units.py filterfx --instances 65536 --frames 12000 --steps 1
cap k_chorus       k_chorus   1 python tools/bench_units.py chorus   --instances 8192  --frames 4800  --steps 1
cap k_bq_tma_fast  k_bq_tma   1 python tools/run_one.py 65536 24000 fast 0 1
cap k_bq_skew_4096 k_bq_skew  1 python tools/run_one.py 4096 48000 fast 0 1
cap k_bq_sys_4096  k_bq_sys   1 python tools/run_one.py 4096 48000 fast 3 1
cap k_bq_skew_8192 k_bq_skew  1 python tools/run_one.py 8192 24000 fast 0 1
# compute-sanitizer is closed on this pool ("runs under it have left GPUs needing a reset", gpurun r2 run 11, rc 86 for every
# tool): the multi-warp mbarrier pipelines are covered by bit-exact parity at many shapes and by the 10 s runs instead.
ls -la $O | head -50
