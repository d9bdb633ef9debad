#!/bin/bash
# N-GPU scaling work: multi-GPU tests, then bench.py --exposed_comm over bucket sizes / all-reduce variants.
#   gpurun --gpus N --timeout 900 -- 'bash tools/calls/r2_scale_sweep.sh N'
set -u
N=${1:-2}
TAG=${2:-s1}
mkdir -p gpurun_out
OUT=gpurun_out/${TAG}_n${N}.jsonl
: > "$OUT"
PORT=29800
python -m huggingface_sagemaker_tensorflow_distributed_b200.build > gpurun_out/${TAG}_build.log 2>&1
if [ "${SKIP_TESTS:-0}" != "1" ]; then
  timeout 600 python -m pytest tests/test_multigpu.py -x -q > gpurun_out/${TAG}_n${N}_tests.log 2>&1
  echo "tests rc=$?" >> gpurun_out/${TAG}_n${N}_tests.log
  tail -5 gpurun_out/${TAG}_n${N}_tests.log
fi
run() {  # label, env assignments..., -- bench flags
  local label=$1; shift
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done
  shift
  PORT=$((PORT + 1))
  env "${envs[@]}" timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 \
    --master-port $PORT bench.py --gpus "$N" "$@" 2> gpurun_out/${TAG}_err.log | grep "^{" | \
    python -c "import sys,json
for l in sys.stdin:
    d=json.loads(l); d['label']='$label'; print(json.dumps(d))" >> "$OUT" || echo "{\"failed\": \"$label\"}" >> "$OUT"
  tail -1 "$OUT" | python -c "import sys,json
d=json.loads(sys.stdin.read())
print('$label', d.get('ms_per_step'), 'ms', d.get('value'), 'samples/s exposed', d.get('exposed_comm_ms_per_step'), 'nocomm', d.get('no_comm_ms_per_step'))" 2>/dev/null || tail -3 gpurun_out/${TAG}_err.log
}
S="--steps 100 --warmup 8 --exposed_comm"
run "mb32" HSTD_X=1 -- $S --bucket_mb 32 --timeline gpurun_out/${TAG}_timeline_n${N}.txt
run "mb32-two_shot" HSTD_AR_MODE=two_shot -- $S --bucket_mb 32
run "mb32-nvls" HSTD_AR_MODE=nvls -- $S --bucket_mb 32
run "mb16" HSTD_X=1 -- $S --bucket_mb 16
if [ "${MORE:-0}" = "1" ]; then
  run "mb32-one-launch" HSTD_AR_SPLIT=0 -- $S --bucket_mb 32
  run "mb32-dense-emb" HSTD_SPARSE_EMB=0 -- $S --bucket_mb 32
fi
