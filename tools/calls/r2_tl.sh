#!/bin/bash
# N-GPU: bench with exposed-comm and a kernel timeline of one step on rank 0
set -u
N=${1:-2}; TAG=${2:-tl}; MB=${3:-16}
mkdir -p gpurun_out
python -m huggingface_sagemaker_tensorflow_distributed_b200.build > gpurun_out/${TAG}_build.log 2>&1
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29911 \
  bench.py --gpus $N --steps 100 --warmup 8 --exposed_comm --bucket_mb $MB --timeline gpurun_out/${TAG}_timeline_n${N}.txt \
  > gpurun_out/${TAG}_n${N}.json 2> gpurun_out/${TAG}_err.log
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_n${N}.json").read().strip().splitlines()[-1])
    print(d["ms_per_step"], "ms", d["value"], "exposed", d.get("exposed_comm_ms_per_step"), "nocomm", d.get("no_comm_ms_per_step"), d.get("dp"))
except Exception as e:
    print("failed", e); print(open("gpurun_out/${TAG}_err.log").read()[-2000:])
PY
head -3 gpurun_out/${TAG}_timeline_n${N}.txt; tail -1 gpurun_out/${TAG}_timeline_n${N}.txt
