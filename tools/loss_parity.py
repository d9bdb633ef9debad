"""Loss-curve parity: this framework vs the PyTorch DDP + cuBLAS/SDPA library arm, same seed, same initial
weights, same synthetic batches in the same order (the History the reference writes at scripts/train.py:159-165
is what gets compared).  Per-step loss of both arms -> JSONL; prints the max / mean |difference| of the raw and
the smoothed (window 10) curves.  Run with python (1 GPU) or torchrun (N GPUs).

    python tools/loss_parity.py --model distilbert-base-uncased --steps 300 --dropout 0 --out profiles/loss.jsonl
"""
import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline"))
import huggingface_sagemaker_tensorflow_distributed_b200.parallel.hvd as hvd  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200 import trainer as T  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.data import BatchedDataset, synthetic_imdb  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.models.configuration import resolve_config  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.optim import Adam, scale_learning_rate  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.parallel import comm  # noqa: E402


def smooth(x, w=10):
    out, acc = [], 0.0
    for i, v in enumerate(x):
        acc += v
        if i >= w:
            acc -= x[i - w]
        out.append(acc / min(i + 1, w))
    return out


def run(model_name="distilbert-base-uncased", steps=300, batch=8, seq=512, dropout=0.0, lr=5e-5, seed=42, out=None,
        log=print):
    import ddp_baseline

    hvd.init()
    rank, world, dev = hvd.rank(), hvd.size(), comm.device()
    cfg = resolve_config(model_name)
    cfg = cfg.replace(hidden_dropout=dropout, attention_dropout=dropout, classifier_dropout=dropout)
    ds = BatchedDataset(synthetic_imdb(cfg, steps * batch * world, seq, seed=1), batch, rank, world, pin_memory=True)
    batches = list(ds)
    lr_w = scale_learning_rate(lr, world)
    # ---- this framework
    model = T.AutoModelForSequenceClassification.from_pretrained(model_name, device=dev, dropout=dropout, seed=seed)
    opt = hvd.DistributedOptimizer(Adam(learning_rate=lr_w))
    model.compile(optimizer=opt, loss=T.SparseCategoricalCrossentropy(from_logits=True))
    eng = model.backend
    eng.setup_training(opt, 10 ** 6, batch, seq)
    model.broadcast_state(0)
    sd0 = {k: v.clone() for k, v in model.state_dict().items()}
    ours = torch.zeros(steps, 3, device=dev)
    for i in range(steps):
        eng.train_step(*batches[i])
        ours[i] = eng.last_step_metrics()[:3]
    torch.cuda.synchronize()
    eng.finish()
    # ---- library arm from the SAME initial weights
    ref, ropt = ddp_baseline.build(cfg, dev, world, lr_w)
    core = ref.module if world > 1 else ref
    core.load_state_dict({k: v.to(dev) for k, v in sd0.items()})
    theirs = torch.zeros(steps, 2, device=dev)
    for i in range(steps):
        f, l = batches[i]
        ids, mask, lab = f["input_ids"].to(dev).long(), f["attention_mask"].to(dev), l.to(dev).long()
        ropt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logits = ref(ids, mask)
        loss = F.cross_entropy(logits.float(), lab, ignore_index=-1)
        loss.backward()
        ropt.step()
        theirs[i, 0], theirs[i, 1] = loss.detach() * (lab >= 0).sum(), (lab >= 0).sum()
    # global mean loss per step (sum over ranks of loss_sum / sum of counts)
    o = torch.stack([ours[:, 0], ours[:, 2]], 1)
    if world > 1:
        import torch.distributed as dist

        dist.all_reduce(o)
        dist.all_reduce(theirs)
    lo = (o[:, 0] / o[:, 1].clamp(min=1)).tolist()
    lt = (theirs[:, 0] / theirs[:, 1].clamp(min=1)).tolist()
    so, st_ = smooth(lo), smooth(lt)
    d_raw = [abs(a - b) for a, b in zip(lo, lt)]
    d_sm = [abs(a - b) for a, b in zip(so, st_)]
    d_w50 = [abs(a - b) for a, b in zip(smooth(lo, 50), smooth(lt, 50))]

    def crossing(x, thr):  # first step whose 10-step mean is below thr
        return next((i + 1 for i, v in enumerate(smooth(x)) if i >= 9 and v < thr), None)

    res = {"model": model_name, "n_gpus": world, "steps": steps, "per_gpu_batch": batch, "seq_len": seq, "dropout": dropout,
           "max_abs_diff": max(d_raw), "mean_abs_diff": sum(d_raw) / len(d_raw), "max_abs_diff_smoothed": max(d_sm),
           # bf16 training is chaotic once the loss starts to collapse: identical for the first tens of steps, then the
           # two arms spike at different steps while following the same envelope (50-step means, threshold crossings)
           "max_abs_diff_first50": max(d_raw[:50]), "max_abs_diff_window50": max(d_w50[49:]) if steps >= 50 else None,
           "steps_to_loss_below": {str(t): [crossing(lo, t), crossing(lt, t)] for t in (0.5, 0.3, 0.1)},
           "first_loss": [lo[0], lt[0]], "last10_mean": [sum(lo[-10:]) / 10, sum(lt[-10:]) / 10]}
    if rank == 0:
        log(json.dumps(res))
        if out:
            os.makedirs(os.path.dirname(os.path.abspath(out)), exist_ok=True)
            with open(out, "a") as f:
                f.write(json.dumps({**res, "summary": True}) + "\n")
                for i, (a, b) in enumerate(zip(lo, lt)):
                    f.write(json.dumps({"model": model_name, "n_gpus": world, "dropout": dropout, "step": i + 1,
                                        "loss_ours": round(a, 6), "loss_ddp": round(b, 6)}) + "\n")
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="distilbert-base-uncased")
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--seq", type=int, default=512)
    ap.add_argument("--dropout", type=float, default=0.0)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    run(a.model, a.steps, a.batch, a.seq, a.dropout, out=a.out)
    hvd.shutdown()
