"""B200 engine vs the PyTorch oracle: one full training step (loss, logits, every gradient),
then short training runs (graph replay == eager, loss goes down, checkpoint round trip)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _mk(name, **kw):
    from huggingface_sagemaker_tensorflow_distributed_b200 import trainer as T
    return T.AutoModelForSequenceClassification.from_pretrained(name, device=torch.device("cuda"), **kw)


def _batch(cfg, B, S, seed=0):
    from huggingface_sagemaker_tensorflow_distributed_b200.data import BatchedDataset, synthetic_imdb
    ds = synthetic_imdb(cfg, B, S, seed=seed)
    ds.attention_mask[:, :] = 1
    ds.attention_mask[0, S // 2:] = 0  # one padded sequence
    ds.input_ids[0, S // 2:] = cfg.pad_token_id
    return next(iter(BatchedDataset(ds, B)))


@pytest.mark.parametrize("name,fused", [("tiny-distilbert-gpu", True), ("tiny-bert-gpu", True),
                                        ("tiny-roberta-gpu", True), ("tiny-bert-gpu", False)])
def test_one_step_matches_autograd_oracle(name, fused):
    from huggingface_sagemaker_tensorflow_distributed_b200.models.reference import EncoderForSequenceClassification
    from huggingface_sagemaker_tensorflow_distributed_b200.optim import Adam
    model = _mk(name, dropout=0.0, use_cuda_graph=False, fused_attention=fused)
    eng, cfg = model.backend, model.cfg
    B, S = 4, 256
    feats, labels = _batch(cfg, B, S)
    model.compile(optimizer=Adam(learning_rate=0.0))
    eng.setup_training(model.optimizer, 10, B, S)
    # oracle with the bf16-rounded weights the engine actually multiplies with
    ref = EncoderForSequenceClassification(cfg).cuda().train()
    ref.load_state_dict({n: eng.W(n).float() for n in eng.slots})
    eng.reset_metrics()
    eng.train_step(feats, labels)
    torch.cuda.synchronize()
    ids, mask, lab = feats["input_ids"].cuda().long(), feats["attention_mask"].cuda(), labels.cuda().long()
    logits = ref(ids, mask)
    loss = F.cross_entropy(logits, lab)
    loss.backward()
    got = eng.logits(B, S, True)
    assert torch.allclose(got, logits, atol=3e-2, rtol=3e-2), (got, logits)
    m = eng.read_metrics()
    assert m["loss"] == pytest.approx(loss.item(), rel=2e-2) and m["count"] == B
    bad = []
    for n, p in ref.named_parameters():
        g, gr = eng.Gr(n).float(), p.grad
        rel = (g - gr).norm() / (gr.norm() + 1e-12)
        if not (rel < 0.09 or (g - gr).abs().max() < 2e-4):
            bad.append((n, rel.item(), gr.norm().item()))
    assert not bad, bad


def _per_step_losses(model, ds, steps):
    out = []
    it = iter(ds)
    for _ in range(steps):
        feats, labels = next(it)
        model.backend.train_step(feats, labels)
        out.append(model.backend.last_step_metrics().clone())
    torch.cuda.synchronize()
    return [(m[0] / m[2]).item() for m in out]


def test_graph_replay_equals_eager():
    from huggingface_sagemaker_tensorflow_distributed_b200.data import BatchedDataset, synthetic_imdb
    from huggingface_sagemaker_tensorflow_distributed_b200.optim import Adam
    runs = []
    for graph in (False, True):
        model = _mk("tiny-distilbert-gpu", dropout=0.0, use_cuda_graph=graph, seed=7)
        ds = BatchedDataset(synthetic_imdb(model.cfg, 64, 128, seed=3), 8)
        model.compile(optimizer=Adam(learning_rate=1e-4))
        model.backend.setup_training(model.optimizer, 8, 8, 128)
        runs.append(_per_step_losses(model, ds, 8))
        assert (model.backend.plans[(8, 128, True)].graph is not None) == graph
    # (bf16 atomics in the embedding scatter make the two runs agree to rounding, not bitwise)
    assert runs[0] == pytest.approx(runs[1], rel=2e-3), runs


def test_training_tracks_the_fp32_oracle_and_learns():
    """Same data, same init, dropout off: per-step loss of the bf16 engine vs fp32 PyTorch."""
    from huggingface_sagemaker_tensorflow_distributed_b200 import trainer as T
    from huggingface_sagemaker_tensorflow_distributed_b200.data import BatchedDataset, synthetic_imdb
    from huggingface_sagemaker_tensorflow_distributed_b200.optim import Adam
    model = _mk("tiny-bert-gpu", dropout=0.0, seed=11)
    ref = _mk("tiny-bert-gpu", backend="torch", dtype="fp32", dropout=0.0, seed=11)
    ds = BatchedDataset(synthetic_imdb(model.cfg, 512, 128, seed=5), 16)
    for mdl in (model, ref):
        mdl.compile(optimizer=Adam(learning_rate=1e-4, weight_decay=0.01, lr_schedule="linear", warmup_steps=2))
        mdl.backend.setup_training(mdl.optimizer, 12, 16, 128)
    eng = _per_step_losses(model, ds, 12)
    want = []
    for (feats, labels), _ in zip(ds, range(12)):
        l, c, n = ref.backend.train_step(feats, labels)
        want.append((l / n).item())
    assert eng == pytest.approx(want, abs=2e-2), (eng, want)
    # and a longer run with dropout on learns the planted signal
    model = _mk("tiny-bert-gpu", seed=11)
    model.compile(optimizer=Adam(learning_rate=3e-4), loss=T.SparseCategoricalCrossentropy(from_logits=True))
    loss = model.fit(ds, epochs=5, verbose=0).history["loss"]
    assert loss[-1] < loss[0] - 0.05, loss


def test_dropout_is_active_in_training_only():
    from huggingface_sagemaker_tensorflow_distributed_b200.optim import Adam
    model = _mk("tiny-distilbert-gpu", seed=3, use_cuda_graph=False)
    eng = model.backend
    feats, labels = _batch(model.cfg, 4, 128)
    model.compile(optimizer=Adam(learning_rate=0.0))
    eng.setup_training(model.optimizer, 4, 4, 128)
    eng.eval_step(feats, labels)
    e1 = eng.logits(4, 128, False).clone()
    eng.eval_step(feats, labels)
    assert torch.equal(e1, eng.logits(4, 128, False))
    eng.train_step(feats, labels)
    t1 = eng.logits(4, 128, True).clone()
    eng.train_step(feats, labels)  # lr = 0: only the dropout masks changed
    t2 = eng.logits(4, 128, True)
    assert not torch.allclose(t1, t2) and not torch.allclose(t1, e1)


def test_eval_and_checkpoint_roundtrip(tmp_path):
    import transformers
    from huggingface_sagemaker_tensorflow_distributed_b200.data import BatchedDataset, synthetic_imdb
    model = _mk("tiny-roberta-gpu")
    ds = BatchedDataset(synthetic_imdb(model.cfg, 30, 128, seed=1), 4)  # short last batch (30 = 7*4 + 2)
    res = model.evaluate(ds, return_dict=True)
    model.save_pretrained(str(tmp_path))
    hf = transformers.RobertaForSequenceClassification.from_pretrained(str(tmp_path)).cuda().eval()
    tot, n = 0.0, 0
    with torch.no_grad():
        for feats, labels in ds:
            lab = labels.cuda().long()
            out = hf(input_ids=feats["input_ids"].cuda().long(), attention_mask=feats["attention_mask"].cuda()).logits
            v = lab >= 0
            tot += F.cross_entropy(out[v], lab[v], reduction="sum").item()
            n += int(v.sum())
    assert n == 30 and res["loss"] == pytest.approx(tot / n, rel=3e-2)
    again = _mk(str(tmp_path))
    assert again.evaluate(ds, return_dict=True)["loss"] == pytest.approx(res["loss"], rel=1e-3)
