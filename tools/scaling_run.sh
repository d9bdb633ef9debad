#!/bin/bash
# ours vs DDP library baseline at N GPUs (+ exposed comm), appended to gpurun_out/scaling.jsonl
N=${1:-8}; MODEL=${2:-distilbert-base-uncased}; STEPS=${3:-30}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
if [ "$N" = "1" ]; then TR="python"; fi
mkdir -p gpurun_out
timeout 300 $TR bench.py --gpus $N --steps $STEPS --warmup 5 --model $MODEL --exposed_comm 2>&1 | grep "^{" | tee -a gpurun_out/scaling.jsonl | cut -c1-180
timeout 300 $TR bench.py --gpus $N --steps $STEPS --warmup 5 --model $MODEL --impl ddp 2>&1 | grep "^{" | tee -a gpurun_out/scaling.jsonl | cut -c1-180
