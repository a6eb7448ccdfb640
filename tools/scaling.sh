#!/bin/bash
# Multi-GPU runs of bench.py on one box (JSON lines appended to gpurun_out/scaling.jsonl).
# Usage: bash tools/scaling.sh "2 4 8"
for n in $1; do
  timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
    --master-port $((29600 + n)) bench.py --gpus $n --steps 3 --warmup 3 >> gpurun_out/scaling.jsonl 2>> gpurun_out/scaling.log
  echo "n=$n rc=$?"
done
