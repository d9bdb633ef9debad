"""The product's entry points on a real GPU (reference launch.py:36-55 -> scripts/train.py:139-183,
scripts/singe_node_train.py:40-120): the same scripts a user runs, as subprocesses, with the SageMaker
environment contract — not the engine poked directly."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _kv(path):
    out = {}
    for line in open(path):
        k, _, v = line.partition(" = ")
        out[k.strip()] = v.strip()
    return out


def _run(cmd, env, timeout=900):
    r = subprocess.run(cmd, env=env, cwd=REPO, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, (r.stdout[-3000:], r.stderr[-3000:])
    return r.stdout + r.stderr


def test_train_py_runs_distilbert_base_for_20_steps(tmp_path):
    """scripts/train.py, world size 1 (the reference's checked-in smoke configuration: distribution=None on a
    one-GPU instance), DistilBERT-base at S=512, per-rank batch 8: fit -> train_results.txt -> evaluate ->
    eval_results.txt -> save_pretrained, then the checkpoint reloads into transformers."""
    env = dict(os.environ, SM_MODEL_DIR=str(tmp_path / "model"), SM_OUTPUT_DATA_DIR=str(tmp_path / "out"),
               SM_NUM_GPUS="1", CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", "0").split(",")[0])
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    log = _run([sys.executable, "scripts/train.py", "--epochs", "1", "--train_batch_size", "8", "--eval_batch_size", "2",
                "--model_name_or_path", "distilbert-base-uncased", "--max_steps", "20", "--max_eval_steps", "10",
                "--train_samples", "160", "--test_samples", "20", "--log_every", "10"], env)
    tr = _kv(tmp_path / "out" / "train_results.txt")
    ev = _kv(tmp_path / "out" / "eval_results.txt")
    loss = json.loads(tr["loss"])
    assert len(loss) == 1 and 0.0 < loss[0] < 2.0, tr
    assert "train_runtime" in tr["train_runtime"] and 0.0 <= float(ev["sparse_categorical_accuracy"]) <= 1.0
    assert 0.0 < float(ev["loss"]) < 5.0
    assert "samples/s" in log
    files = set(os.listdir(tmp_path / "model"))
    assert {"config.json", "model.safetensors"} <= files, files
    from transformers import AutoModelForSequenceClassification

    m = AutoModelForSequenceClassification.from_pretrained(str(tmp_path / "model"))
    assert sum(p.numel() for p in m.parameters()) == 66_955_010


def test_singe_node_train_and_launcher(tmp_path):
    """The single-process script on this box's GPUs (MirroredStrategy front-end; with several GPUs it becomes one
    replica process per GPU), then launch.py -> LocalEstimator.fit() -> scripts/train.py with the artifacts
    tarred the way SageMaker leaves them."""
    env = dict(os.environ, SM_MODEL_DIR=str(tmp_path / "m1"), SM_OUTPUT_DATA_DIR=str(tmp_path / "o1"))
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    _run([sys.executable, "scripts/singe_node_train.py", "--epochs", "1", "--train_batch_size", "16", "--eval_batch_size", "4",
          "--model_name_or_path", "tiny-distilbert-gpu", "--max_seq_length", "128", "--train_samples", "256",
          "--test_samples", "32"], env)
    tr = _kv(tmp_path / "o1" / "train_results.txt")
    assert "train_runtime" not in tr and len(json.loads(tr["loss"])) == 1  # (the reference's single-node script records none)
    assert os.path.exists(tmp_path / "m1" / "model.safetensors")
    out = _run([sys.executable, "launch.py", "--distribution", "mpi", "--instance_type", "local_gpu", "--model",
                "tiny-distilbert-gpu", "--output_path", str(tmp_path / "jobs"), "--hyperparameters",
                json.dumps({"max_seq_length": 128, "train_samples": 256, "test_samples": 32})], env)
    job = json.loads(out[out.index("{\n"):out.rindex("}") + 1])
    assert os.path.exists(job["model_data"]) and os.path.exists(job["output_data"])
    assert os.path.exists(os.path.join(job["output_data_dir"], "eval_results.txt"))
