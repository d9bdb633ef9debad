#!/bin/bash
# One-call-per-GPU-count measurement plan for the BASELINE.json configs that still lack numbers
# (BERT-base at 1/2/4/8, RoBERTa-large at 8, all-reduce sweeps at 2/4) plus a refresh of the headline
# at 8 GPUs with the final kernels.  Every command runs under `timeout`; outputs go to distinct
# gpurun_out/ files (the box starts with an empty gpurun_out/, merged files overwrite same names).
#
#   gpurun --gpus 8 --timeout 420 -- 'bash tools/round2_gpu_plan.sh 8'     (≈ 4 min x 8 GPUs of budget)
#   gpurun --gpus 4 --timeout 200 -- 'bash tools/round2_gpu_plan.sh 4'
#   gpurun --gpus 2 --timeout 200 -- 'bash tools/round2_gpu_plan.sh 2'
set -u
N=${1:-8}
OUT=gpurun_out/plan_n${N}.jsonl
mkdir -p gpurun_out
: > "$OUT"
PORT=29700
run() {  # model steps warmup [extra bench.py flags...]
  local model=$1 steps=$2 warm=$3; shift 3
  PORT=$((PORT + 1))
  local tr="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT"
  [ "$N" = "1" ] && tr="python"
  timeout 90 $tr bench.py --gpus "$N" --steps "$steps" --warmup "$warm" --model "$model" "$@" 2>gpurun_out/plan_err.log | grep "^{" >> "$OUT" \
    || echo "{\"failed\": \"$model $*\", \"n_gpus\": $N}" >> "$OUT"
}
run distilbert-base-uncased 60 8 --exposed_comm
run bert-base-uncased 40 6 --exposed_comm
run bert-large-uncased 20 4
if [ "$N" = "8" ]; then
  run roberta-large 20 4
  run bert-base-uncased 30 5 --impl ddp
fi
if [ "$N" != "1" ]; then
  PORT=$((PORT + 1))
  SWEEP_MAX_BYTES=$((1 << 28)) timeout 90 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 \
    --master-port $PORT tools/allreduce_sweep.py > gpurun_out/plan_sweep_n${N}.log 2>&1 || echo "sweep failed" >> gpurun_out/plan_sweep_n${N}.log
fi
python - "$OUT" <<'PY'
import json, sys
for line in open(sys.argv[1]):
    try:
        d = json.loads(line)
    except ValueError:
        continue
    if "failed" in d:
        print("FAILED", d["failed"])
        continue
    print(f'{d["config"]["model"]:28s} n={d["n_gpus"]} {d.get("impl","ours"):5s} {d["ms_per_step"]:8.3f} ms {d["value"]:10.1f} samples/s'
          f' e2e {d["e2e"]["value"]:10.1f} exposed {d.get("exposed_comm_ms_per_step")}')
PY
