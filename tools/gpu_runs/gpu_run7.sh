#!/bin/bash
# 8-GPU box: headline bench at N = 1, 2, 4, 8 (weak scaling, no collective) and the voice bank with its NCCL bus reduce
set -u
O=gpurun_out; mkdir -p $O
nvidia-smi -L | head -8
python bench.py --steps 100 --no-cpu-baseline > $O/scale_n1.json 2> $O/scale.err; cut -c1-200 $O/scale_n1.json
p=29500
for n in 2 4 8; do
  p=$((p+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $p bench.py --gpus $n --steps 100 > $O/scale_n$n.json 2>> $O/scale.err
  cut -c1-200 $O/scale_n$n.json
done
for n in 2 8; do
  p=$((p+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $p tools/bench_units.py voices --instances 65536 --frames 4800 --steps 10 > $O/voices_n$n.json 2>> $O/scale.err
  cut -c1-400 $O/voices_n$n.json
done
tail -3 $O/scale.err
