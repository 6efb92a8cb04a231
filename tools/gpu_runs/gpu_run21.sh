#!/bin/bash
# 2-GPU check of the final tree: headline bench under torchrun, the voice bank with its NCCL bus reduce
set -u
O=gpurun_out; mkdir -p $O
nvidia-smi -L | head -4
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 > $O/final_n2.json 2> $O/final_n2.err; echo "bench rc=$?"; cut -c1-260 $O/final_n2.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/bench_units.py voices --instances 65536 --frames 4800 --steps 10 > $O/final_voices_n2.json 2>> $O/final_n2.err; echo "voices rc=$?"; cut -c1-300 $O/final_voices_n2.json
timeout 300 python tools/bench_units.py voices --instances 65536 --frames 4800 --steps 10 2>/dev/null | grep '^{' > $O/final_voices_n1.json; cut -c1-200 $O/final_voices_n1.json
tail -3 $O/final_n2.err
