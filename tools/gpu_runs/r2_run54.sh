#!/bin/bash
# round 2, run 54 (8-GPU box): bench.py at N = 1, 4, 8 with k_bq_roll in the strong-scaling shards (16 384 and 8 192 channels per GPU)
set -u
O=gpurun_out/r2_run54; mkdir -p $O
nvidia-smi -L | head -8
python bench.py --steps 30 --no-cpu-baseline > $O/bench_n1.json 2> $O/bench.err; echo "n1 rc=$?"
p=29700
for n in 4 8; do
  p=$((p+1))
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $p bench.py --gpus $n --steps 30 > $O/bench_n$n.json 2>> $O/bench.err; echo "n$n rc=$?"
done
python - <<'PY'
import json
base=None
for n in (1,4,8):
    try: d=json.load(open(f'gpurun_out/r2_run54/bench_n{n}.json'))
    except Exception as e: print(n, 'missing', e); continue
    if n==1: base=d
    s=d.get('strong',{}); v=d.get('voices',{}); e=d.get('e2e',{})
    print(f"N={n} value {d['value']:.4e} ({d['value']/base['value']:.2f}x) strong {s.get('value',0):.4e} ({s.get('value',0)/base['value']:.2f}x, {s.get('kernel')}, frac {s.get('roofline_frac',0):.3f}) "
          f"voices {v.get('value',0):.4e} ({v.get('value',0)/base['voices']['value']:.2f}x) e2e {e.get('value',0):.3e} pcie_frac {e.get('pcie_frac')}")
PY
tail -3 $O/bench.err
