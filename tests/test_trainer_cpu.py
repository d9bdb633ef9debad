"""Training runtime on CPU: Keras-shaped fit/evaluate, results files, DP equivalence over
gloo (unit + multi-rank-without-GPUs tiers)."""
import os
import re
import subprocess
import sys

import pytest
import torch

from huggingface_sagemaker_tensorflow_distributed_b200 import trainer as T
from huggingface_sagemaker_tensorflow_distributed_b200.data import BatchedDataset, synthetic_imdb
from huggingface_sagemaker_tensorflow_distributed_b200.optim import Adam

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _tiny(seed=0, dropout=0.0):
    return T.AutoModelForSequenceClassification.from_pretrained("tiny-distilbert", device=torch.device("cpu"),
                                                                seed=seed, dropout=dropout)


def test_fit_history_evaluate_and_results_files(tmp_path):
    torch.manual_seed(0)
    model = _tiny()
    ds = BatchedDataset(synthetic_imdb(model.cfg, 96, 32), 8)
    model.compile(optimizer=Adam(learning_rate=2e-3), loss=T.SparseCategoricalCrossentropy(from_logits=True),
                  metrics=[T.SparseCategoricalAccuracy()])
    hist = model.fit(ds, epochs=3, verbose=0)
    assert sorted(hist.history) == ["loss", "sparse_categorical_accuracy"]
    assert len(hist.history["loss"]) == 3 and hist.history["loss"][-1] < hist.history["loss"][0]
    res = model.evaluate(ds, return_dict=True)
    assert set(res) == {"loss", "sparse_categorical_accuracy"}
    T.write_train_results(str(tmp_path / "train_results.txt"), hist, 12.34567)
    T.write_eval_results(str(tmp_path / "eval_results.txt"), res)
    lines = open(tmp_path / "train_results.txt").read().splitlines()
    assert re.fullmatch(r"loss = \[[0-9.e\-, ]+\]", lines[0])
    assert lines[-1] == "train_runtime = {'train_runtime': 12.3457}"  # scripts/train.py:154,165
    assert open(tmp_path / "eval_results.txt").read().startswith("loss = ")


def test_compile_rejects_unsupported_loss():
    with pytest.raises(ValueError):
        T.SparseCategoricalCrossentropy(from_logits=False)


def test_short_last_batch_is_ignored_in_loss_and_accuracy():
    model = _tiny()
    ds_full = synthetic_imdb(model.cfg, 10, 16)
    a = model.evaluate(BatchedDataset(ds_full, 4), return_dict=True)  # 4+4+2(padded)
    b = model.evaluate(BatchedDataset(ds_full, 10), return_dict=True)
    assert a["loss"] == pytest.approx(b["loss"], rel=1e-5)
    assert a["sparse_categorical_accuracy"] == pytest.approx(b["sparse_categorical_accuracy"])


DP_SCRIPT = r"""
import os, sys, torch
sys.path.insert(0, {repo!r})
os.environ["HSTD_FORCE_CPU"] = "1"
import huggingface_sagemaker_tensorflow_distributed_b200.parallel.hvd as hvd
from huggingface_sagemaker_tensorflow_distributed_b200 import trainer as T
from huggingface_sagemaker_tensorflow_distributed_b200.data import BatchedDataset, synthetic_imdb
from huggingface_sagemaker_tensorflow_distributed_b200.optim import Adam
hvd.init()
torch.set_num_threads(1)
model = T.AutoModelForSequenceClassification.from_pretrained("tiny-distilbert", device=torch.device("cpu"),
                                                            seed=100 + hvd.rank(), dropout=0.0)  # different init per rank
ds = synthetic_imdb(model.cfg, 32, 16)
opt = hvd.DistributedOptimizer(Adam(learning_rate=1e-3))
model.compile(optimizer=opt, loss=T.SparseCategoricalCrossentropy(from_logits=True))
model.fit(BatchedDataset(ds, 4, hvd.rank(), hvd.size()), epochs=1, verbose=0,
          callbacks=[hvd.callbacks.BroadcastGlobalVariablesCallback(0)])
model.save_pretrained({out!r} + f"/w{{hvd.size()}}")
sd = model.state_dict()
torch.save(sd, {out!r} + f"/sd_w{{hvd.size()}}_r{{hvd.rank()}}.pt")
"""


@pytest.mark.timeout(300)
def test_two_rank_gloo_equals_one_rank_with_double_batch(tmp_path):
    """2 ranks × bs 4 (sharded, grads averaged) ≡ 1 rank × bs 8 — and the rank-0 broadcast
    makes differently-initialised ranks agree (scripts/train.py:112-114,133)."""
    script = tmp_path / "dp.py"
    script.write_text(DP_SCRIPT.format(repo=REPO, out=str(tmp_path)))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29541", WORLD_SIZE="2", OMP_NUM_THREADS="1")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r), LOCAL_RANK=str(r)))
             for r in range(2)]
    assert [p.wait(timeout=240) for p in procs] == [0, 0]
    sd0 = torch.load(tmp_path / "sd_w2_r0.pt")
    sd1 = torch.load(tmp_path / "sd_w2_r1.pt")
    for k in sd0:
        assert torch.equal(sd0[k], sd1[k]), f"ranks diverged on {k}"
    assert sorted(os.listdir(tmp_path / "w2")) == ["config.json", "model.safetensors", "special_tokens_map.json",
                                                   "tokenizer_config.json", "vocab.txt"]
    # single process, same init as rank 0, same samples per step in one batch of 8
    model = T.AutoModelForSequenceClassification.from_pretrained("tiny-distilbert", device=torch.device("cpu"),
                                                                seed=100, dropout=0.0)
    ds = synthetic_imdb(model.cfg, 32, 16)

    class Interleaved(BatchedDataset):  # batch b = samples {8b..8b+7}: what ranks 0,1 saw together
        pass

    model.compile(optimizer=Adam(learning_rate=1e-3), loss=T.SparseCategoricalCrossentropy(from_logits=True))
    model.fit(Interleaved(ds, 8), epochs=1, verbose=0)
    ref = model.state_dict()
    for k in ref:
        assert torch.allclose(ref[k], sd0[k], atol=2e-5), k


def test_true_resume_with_optimizer_state(tmp_path):
    """2 steps + save(training state) + reload + 2 steps == 4 steps straight (beyond the
    reference, whose only resume is weights via --model_name_or_path)."""
    ds = None
    def run(model, batches):
        model.compile(optimizer=Adam(learning_rate=1e-3))
        model.backend.setup_training(model.optimizer, 100, 4, 16)
        for feats, labels in batches:
            model.backend.train_step(feats, labels)
    base = _tiny(seed=3)
    ds = list(BatchedDataset(synthetic_imdb(base.cfg, 16, 16), 4))
    run(base, ds)
    straight = base.state_dict()
    first = _tiny(seed=3)
    run(first, ds[:2])
    first.save_pretrained(str(tmp_path), save_training_state=True)
    assert "training_state.pt" in os.listdir(tmp_path)
    resumed = T.AutoModelForSequenceClassification.from_pretrained(str(tmp_path), device=torch.device("cpu"), dropout=0.0)
    run(resumed, ds[2:])
    for k, v in straight.items():
        assert torch.allclose(resumed.state_dict()[k], v, atol=1e-6), k
