#!/bin/bash
# round 2, run 11: evidence at HEAD -- ncu --set full of the kernels VERDICT item 6 names, then compute-sanitizer on small shapes
set -u
O=gpurun_out/r2_run11; mkdir -p $O
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
cap k_bq_sys_4096  k_bq_sys   1 python tools/run_one.py 4096 48000 fast 0 1
cap k_bq_skew_8192 k_bq_skew  1 python tools/run_one.py 8192 24000 fast 0 1
# compute-sanitizer: small shapes of the kernels with hand-rolled mbarrier pipelines
san() {  # san <name> -- command...
  n=$1; shift
  for tool in memcheck racecheck; do
    timeout 420 compute-sanitizer --tool $tool --print-limit 5 "$@" > $O/san_${n}_$tool.log 2>&1
    echo "san $n $tool rc=$? :: $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $O/san_${n}_$tool.log | tail -1)"
    [ -z "${SAN_SHOWN:-}" ] && { SAN_SHOWN=1; head -c 1500 $O/san_${n}_$tool.log; }
  done
}
san bq_sys    python tools/run_one.py 1024 256 fast 0 1
san bq_skew   python tools/run_one.py 8192 256 fast 4 1
san fx_g1     env OLB_FX_SPLIT_G=1 python tools/bench_units.py fxrack  --instances 256 --frames 256 --steps 1
san fx_g2     env OLB_FX_SPLIT_G=2 python tools/bench_units.py delayfx --instances 256 --frames 256 --steps 1
san fx_g4     env OLB_FX_SPLIT_G=4 python tools/bench_units.py fxrack  --instances 256 --frames 256 --steps 1
san plate_split env OLB_PLATE_KERNEL=split python tools/bench_units.py dattorro --instances 64 --frames 256 --steps 1
san plate_smem  env OLB_PLATE_KERNEL=smem  python tools/bench_units.py dattorro --instances 8 --frames 256 --steps 1
san chorus    python tools/bench_units.py chorus --instances 64 --frames 256 --steps 1
ls -la $O | head -50
