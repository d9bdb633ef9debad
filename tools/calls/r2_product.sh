#!/bin/bash
# The product on N GPUs: launch.py -> LocalEstimator.fit -> scripts/train.py (full IMDB-sized epoch, sharded and
# reference-faithful unsharded), then scripts/singe_node_train.py; result files copied to gpurun_out/e2e_r2/.
#   gpurun --gpus 8 --timeout 900 -- 'bash tools/calls/r2_product.sh 8'
set -u
N=${1:-8}
OUT=gpurun_out/e2e_r2
mkdir -p $OUT
python -m huggingface_sagemaker_tensorflow_distributed_b200.build > $OUT/build.log 2>&1
job() {  # tag, model, extra hyperparameters json, entry
  local tag=$1 model=$2 hp=$3 entry=${4:-train.py}
  local t0=$(date +%s.%N)
  timeout 600 python launch.py --entry_point $entry --distribution mpi --instance_type local_gpu$N --model $model \
    --output_path /tmp/sm_jobs_$tag --hyperparameters "$hp" > $OUT/${tag}_n${N}.log 2>&1
  local rc=$?
  local t1=$(date +%s.%N)
  local d=$(ls -d /tmp/sm_jobs_$tag/*/ 2>/dev/null | head -1)
  mkdir -p $OUT/${tag}_n${N}
  cp $d/output/data/*.txt $OUT/${tag}_n${N}/ 2>/dev/null
  ls -la $d/model > $OUT/${tag}_n${N}/model_dir_listing.txt 2>/dev/null
  echo "{\"tag\": \"$tag\", \"model\": \"$model\", \"n_gpus\": $N, \"rc\": $rc, \"job_wall_s\": $(python -c "print(round($t1-$t0,2))"), \"entry\": \"$entry\", \"hyperparameters\": $hp}" >> $OUT/jobs_n${N}.jsonl
  grep -E "train_runtime|samples/s \(global\)|Epoch 1/1 -|eval" $OUT/${tag}_n${N}.log | tail -4
  # reload the checkpoint into transformers (rank-0 artefact)
  python - "$d/model" >> $OUT/${tag}_n${N}/reload.txt 2>&1 <<'PY'
import sys
from transformers import AutoModelForSequenceClassification
m = AutoModelForSequenceClassification.from_pretrained(sys.argv[1])
print("reloaded into transformers:", type(m).__name__, sum(p.numel() for p in m.parameters()), "parameters")
PY
  tail -1 $OUT/${tag}_n${N}/reload.txt
}
job distilbert_sharded distilbert-base-uncased '{"eval_batch_size": 64, "log_every": 100}'
if [ "${NO_SHARD:-1}" = "1" ]; then
job distilbert_no_shard distilbert-base-uncased '{"no_shard": true, "eval_batch_size": 64, "max_eval_steps": 50, "log_every": 500}'
fi
job bertlarge_sharded bert-large-uncased-whole-word-masking '{"eval_batch_size": 64, "log_every": 100}'
job distilbert_single_node distilbert-base-uncased "{\"train_batch_size\": $((8*N)), \"eval_batch_size\": 64, \"max_eval_steps\": 50, \"log_every\": 100}" singe_node_train.py
cat $OUT/jobs_n${N}.jsonl
