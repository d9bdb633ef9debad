#!/bin/bash
# Evidence run on N GPUs of one box: collective tests at world N, the BASELINE.json matrix cells for N (ours,
# CUDA-graph DDP library arm), all-reduce sweep vs NCCL with value checks, loss parity, and (PRODUCT=1) the product
# through launch.py.  Everything under `timeout`; results in gpurun_out/ev_n${N}/.
#   gpurun --gpus 8 --timeout 1200 -- 'PRODUCT=1 LARGE=1 bash tools/calls/r2_evidence.sh 8'
set -u
N=${1:-2}
OUT=gpurun_out/ev_n${N}
mkdir -p $OUT
J=$OUT/bench.jsonl
: > $J
PORT=29600
python -m huggingface_sagemaker_tensorflow_distributed_b200.build > $OUT/build.log 2>&1
tr() { PORT=$((PORT + 1)); if [ "$N" = "1" ]; then echo "python"; else echo "python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT"; fi; }
bench() {  # label + bench.py flags
  local label=$1; shift
  timeout ${BENCH_TIMEOUT:-150} $(tr) bench.py --gpus $N "$@" 2> $OUT/err_$label.log | grep "^{" | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); d['label'] = '$label'; print(json.dumps(d))" >> $J
  tail -1 $J | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('$label', d.get('ms_per_step'), 'ms', d.get('value'), 'samples/s  e2e', d.get('e2e', {}).get('value'), 'exposed', d.get('exposed_comm_ms_per_step'), d.get('cuda_graph'))" 2>/dev/null || { echo "FAILED $label"; tail -3 $OUT/err_$label.log; }
}
if [ "$N" != "1" ] && [ "${SKIP_TESTS:-0}" != "1" ]; then
  timeout 500 python -m pytest tests/test_multigpu.py -q -k "${N}]" > $OUT/tests.log 2>&1; echo "tests rc=$?" >> $OUT/tests.log; tail -4 $OUT/tests.log
fi
if [ "${VARIANTS:-0}" = "1" ] && [ "$N" != "1" ]; then
  # schedule x reduction variant at this world size; the fastest becomes the setting for the rest of the call
  for sp in 1 0; do for md in nvls two_shot; do
    HSTD_AR_SPLIT=$sp HSTD_AR_MODE=$md bench "variant-split${sp}-${md}" --steps 100 --warmup 8 --exposed_comm
  done; done
  eval $(python - $J <<'PY'
import json, sys
best = None
for l in open(sys.argv[1]):
    d = json.loads(l)
    if d.get("label", "").startswith("variant-") and (best is None or d["ms_per_step"] < best["ms_per_step"]):
        best = d
if best:
    _, sp, md = best["label"].split("-")
    print(f"export HSTD_AR_SPLIT={sp[-1]} HSTD_AR_MODE={md}")
PY
)
  echo "chosen: HSTD_AR_SPLIT=${HSTD_AR_SPLIT:-} HSTD_AR_MODE=${HSTD_AR_MODE:-}" | tee $OUT/chosen_variant.txt
fi
bench distilbert --steps 200 --warmup 8 --exposed_comm
bench bert-base --model bert-base-uncased --steps 100 --warmup 6 --exposed_comm
if [ "${LARGE:-0}" = "1" ]; then
  bench bert-large --model bert-large-uncased --steps 40 --warmup 4
  bench roberta-large --model roberta-large --steps 40 --warmup 4
fi
BENCH_TIMEOUT=100 bench distilbert-ddp-graph --impl ddp --graph --steps 100 --warmup 5
if [ "${DDP_MORE:-0}" = "1" ]; then
  bench distilbert-ddp-eager --impl ddp --steps 60 --warmup 5
  bench bert-base-ddp-graph --impl ddp --graph --model bert-base-uncased --steps 60 --warmup 5
fi
if [ "$N" != "1" ]; then
  SWEEP_MAX_BYTES=$((1 << ${SWEEP_LOG2:-28})) timeout 240 $(tr) tools/allreduce_sweep.py > $OUT/allreduce_sweep.txt 2> $OUT/err_sweep.log || echo "sweep failed: $(tail -2 $OUT/err_sweep.log)"
  cp gpurun_out/allreduce_sweep_${N}gpu.json $OUT/ 2>/dev/null
  tail -4 $OUT/allreduce_sweep.txt
fi
timeout 300 $(tr) tools/loss_parity.py --model distilbert-base-uncased --steps ${PARITY_STEPS:-300} --dropout 0 --out $OUT/loss_parity.jsonl 2> $OUT/err_parity.log | tail -1
if [ "${PARITY_DROPOUT:-1}" = "1" ]; then
  timeout 300 $(tr) tools/loss_parity.py --model distilbert-base-uncased --steps ${PARITY_STEPS:-300} --dropout 0.1 --out $OUT/loss_parity.jsonl 2>> $OUT/err_parity.log | tail -1
fi
if [ "${LARGE:-0}" = "1" ]; then
  timeout 300 $(tr) tools/loss_parity.py --model bert-large-uncased --steps 100 --dropout 0 --out $OUT/loss_parity.jsonl 2>> $OUT/err_parity.log | tail -1
fi
if [ "${PRODUCT:-0}" = "1" ]; then
  bash tools/calls/r2_product.sh $N 2>&1 | tail -14
fi
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv > $OUT/nvidia_smi_after.csv 2>&1
echo done
