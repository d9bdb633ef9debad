"""Keras-shaped training runtime (SURVEY.md §1 layer G, §2.1 C15-C19).

``SequenceClassifier`` offers the slice of the Keras / HF-TF model API the reference uses:

    model = AutoModelForSequenceClassification.from_pretrained(name)        # train.py:117
    model.compile(optimizer=opt, loss=SparseCategoricalCrossentropy(from_logits=True),
                  metrics=[SparseCategoricalAccuracy()])                     # train.py:118-123
    history = model.fit(train_ds, epochs=..., callbacks=[...], verbose=...)  # train.py:145-153
    result = model.evaluate(test_ds, return_dict=True)                       # train.py:170
    model.save_pretrained(model_dir)                                         # train.py:182

Underneath, a backend executes the steps: the B200 engine (hand-written sm_100a kernels,
CUDA graph, fused P2P all-reduce + AdamW) on GPUs, plain PyTorch on CPU.  Device-side
metrics are accumulated without host syncs and read back once per log interval / epoch.
"""
from __future__ import annotations

import json
import logging
import os
import sys
import time
from typing import Any, Dict, Iterable, List, Optional

import torch

from . import checkpoint
from .data import BatchedDataset
from .models.configuration import EncoderConfig, resolve_config
from .utils.timers import nvtx_range
from .optim import Adam
from .parallel import comm

logger = logging.getLogger(__name__)


class SparseCategoricalCrossentropy:
    """Mean over the batch of −log softmax(logits)[label] (scripts/train.py:118)."""
    name = "loss"

    def __init__(self, from_logits: bool = True):
        if not from_logits:
            raise ValueError("only from_logits=True is supported (the reference's setting)")
        self.from_logits = True


class SparseCategoricalAccuracy:
    """argmax(logits) == label, averaged (scripts/train.py:119)."""
    name = "sparse_categorical_accuracy"


class History:
    """``tf.keras.callbacks.History`` stand-in: ``.history[key]`` is a list, one float per epoch."""

    def __init__(self):
        self.history: Dict[str, List[float]] = {}
        self.epoch: List[int] = []
        self.params: Dict[str, Any] = {}

    def _append(self, epoch: int, logs: Dict[str, float]) -> None:
        self.epoch.append(epoch)
        for k, v in logs.items():
            self.history.setdefault(k, []).append(v)

    def __repr__(self) -> str:
        return f"<History epochs={len(self.epoch)} keys={sorted(self.history)}>"


class _Meter:
    """Running (loss_sum, correct, count) on the device; one host read per ``read()``.
    The engine backend accumulates inside its own kernels (``train_step`` returns None)."""

    def __init__(self, device: torch.device, backend=None):
        self.backend = backend if hasattr(backend, "read_metrics") else None
        if self.backend is not None:
            self.backend.reset_metrics()
        else:
            self.acc = torch.zeros(3, dtype=torch.float64, device=device)

    def update(self, res) -> None:
        if res is None:
            return
        loss_sum, correct, count = res
        self.acc[0] += loss_sum.to(self.acc.dtype) if torch.is_tensor(loss_sum) else loss_sum
        self.acc[1] += correct.to(self.acc.dtype) if torch.is_tensor(correct) else correct
        self.acc[2] += count.to(self.acc.dtype) if torch.is_tensor(count) else count

    def read(self) -> Dict[str, float]:
        if self.backend is not None:
            return self.backend.read_metrics()
        l, c, n = self.acc.tolist()
        n = max(n, 1.0)
        return {"loss": l / n, "sparse_categorical_accuracy": c / n, "count": n}


class SequenceClassifier:
    def __init__(self, cfg: EncoderConfig, backend, tokenizer=None):
        self.cfg = cfg
        self.config = cfg
        self.backend = backend
        self.tokenizer = tokenizer
        self.optimizer: Optional[Adam] = None
        self.metrics_names = ["loss", "sparse_categorical_accuracy"]
        self._compiled = False
        self._metrics_sink = None
        self.last_fit_stats: Dict[str, float] = {}

    # ------------------------------------------------------------------ Keras surface
    def compile(self, optimizer: Adam, loss=None, metrics: Optional[Iterable] = None,
                experimental_run_tf_function: Optional[bool] = None, **_ignored) -> None:
        # experimental_run_tf_function=False is how the reference forces Keras onto the code
        # path that honours hvd.DistributedOptimizer (scripts/train.py:121-123); here the
        # distributed flag on the optimizer is always honoured.
        if loss is not None and not isinstance(loss, SparseCategoricalCrossentropy):
            raise ValueError("loss must be SparseCategoricalCrossentropy(from_logits=True)")
        for m in metrics or []:
            if not isinstance(m, SparseCategoricalAccuracy):
                raise ValueError("only SparseCategoricalAccuracy is supported")
        self.optimizer = optimizer
        self._compiled = True

    def fit(self, dataset: BatchedDataset, epochs: int = 1, callbacks: Optional[list] = None, verbose: int = 1,
            steps_per_epoch: Optional[int] = None, batch_size: Optional[int] = None,
            validation_batch_size: Optional[int] = None, log_every: int = 50, **_ignored) -> History:
        assert self._compiled, "call compile() before fit()"
        callbacks = list(callbacks or [])
        steps = len(dataset) if steps_per_epoch is None else min(len(dataset), steps_per_epoch)
        total_steps = steps * epochs
        self.backend.setup_training(self.optimizer, total_steps, dataset.batch_size, dataset.ds.seq_len)
        for cb in callbacks:
            if hasattr(cb, "on_train_begin"):
                cb.on_train_begin(self)
        hist = History()
        hist.params = {"epochs": epochs, "steps": steps, "verbose": verbose}
        device = self.backend.device
        t_fit = time.time()
        samples_seen = 0
        health_every = max(1, min(log_every, 50)) if log_every > 0 else 50
        for epoch in range(epochs):
            meter = _Meter(device, self.backend)
            t0 = time.time()
            for step, (feats, labels) in enumerate(dataset):
                if step >= steps:
                    break
                with nvtx_range("train_step"):
                    meter.update(self.backend.train_step(feats, labels))
                if (step + 1) % health_every == 0 and hasattr(self.backend, "poll_health"):
                    self.backend.poll_health()  # a dead peer aborts the job now, not at the end of fit
                if verbose and log_every > 0 and (step + 1) % log_every == 0:
                    r = meter.read()
                    dt = time.time() - t0
                    msg = (f"Epoch {epoch + 1}/{epochs} - {step + 1}/{steps} - {dt:.0f}s - loss: {r['loss']:.4f} - "
                           f"sparse_categorical_accuracy: {r['sparse_categorical_accuracy']:.4f} - "
                           f"{r['count'] * comm.size() / max(dt, 1e-9):.1f} samples/s")
                    print(msg, flush=True)
                    self._sink({"event": "train_progress", "epoch": epoch + 1, "step": step + 1, **r})
            r = meter.read()
            if hasattr(self.backend, "finish"):
                self.backend.finish()  # epoch boundary: blocking read of the watchdog word
            samples_seen += int(r.pop("count"))
            logs = dict(r)
            for cb in callbacks:
                if hasattr(cb, "on_epoch_end"):
                    cb.on_epoch_end(self, logs)
            hist._append(epoch, logs)
            if verbose:
                print(f"Epoch {epoch + 1}/{epochs} - {time.time() - t0:.0f}s - loss: {logs['loss']:.4f} - "
                      f"sparse_categorical_accuracy: {logs['sparse_categorical_accuracy']:.4f}", flush=True)
            self._sink({"event": "epoch_end", "epoch": epoch + 1, **logs})
        self.backend.finish()
        wall = time.time() - t_fit
        self.last_fit_stats = {"train_runtime": wall, "samples_per_rank": samples_seen,
                               "train_samples_per_second": samples_seen * comm.size() / max(wall, 1e-9)}
        return hist

    def evaluate(self, dataset: BatchedDataset, batch_size: Optional[int] = None, return_dict: bool = False,
                 verbose: int = 0, steps: Optional[int] = None, **_ignored):
        meter = _Meter(self.backend.device, self.backend)
        n = len(dataset) if steps is None else min(steps, len(dataset))
        for step, (feats, labels) in enumerate(dataset):
            if step >= n:
                break
            with nvtx_range("eval_step"):
                meter.update(self.backend.eval_step(feats, labels))
        r = meter.read()
        r.pop("count")
        self._sink({"event": "evaluate", **r})
        return r if return_dict else [r["loss"], r["sparse_categorical_accuracy"]]

    # ------------------------------------------------------------------ state
    def broadcast_state(self, root: int = 0) -> None:
        self.backend.broadcast_state(root)

    def state_dict(self) -> Dict[str, torch.Tensor]:
        return self.backend.state_dict()

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        self.backend.load_state_dict(sd)

    def save_pretrained(self, save_dir: str, save_training_state: bool = False) -> None:
        """Rank 0 writes, then everyone meets at a barrier (the reference lets every rank
        write the same directory, scripts/train.py:181-183 — a race we do not keep)."""
        sd = self.state_dict()  # collective on the engine backend (gathers sharded state)
        opt_state = self.backend.optimizer_state() if save_training_state else None
        if comm.rank() == 0:
            checkpoint.save_pretrained(save_dir, self.cfg, sd, tokenizer=self.tokenizer)
            if opt_state is not None:
                checkpoint.save_training_state(save_dir, opt_state, step=opt_state.get("iterations", 0), epoch=0)
        comm.barrier()

    def num_parameters(self) -> int:
        return sum(v.numel() for v in self.state_dict().values())

    # ------------------------------------------------------------------ observability
    def set_metrics_sink(self, path: Optional[str]) -> None:
        if path and comm.rank() == 0:
            os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
            self._metrics_sink = open(path, "a")

    def _sink(self, rec: Dict[str, Any]) -> None:
        if self._metrics_sink is not None:
            rec = {"ts": time.time(), **rec}
            self._metrics_sink.write(json.dumps(rec) + "\n")
            self._metrics_sink.flush()


class AutoModelForSequenceClassification:
    """``TFAutoModelForSequenceClassification`` stand-in (scripts/train.py:117)."""

    @staticmethod
    def from_pretrained(name_or_path: str, *, num_labels: Optional[int] = None, device: Optional[torch.device] = None,
                        dtype: str = "auto", backend: str = "auto", seed: int = 42, dropout: Optional[float] = None,
                        **engine_kw) -> SequenceClassifier:
        cfg = resolve_config(name_or_path, num_labels)
        if dropout is not None:
            cfg = cfg.replace(hidden_dropout=dropout, attention_dropout=dropout, classifier_dropout=dropout)
        device = device or comm.device()
        if backend == "auto":
            backend = "engine" if device.type == "cuda" else "torch"
        if dtype == "auto":
            dtype = "bf16" if device.type == "cuda" else "fp32"
        if backend == "engine":
            from .models.engine import EngineBackend

            be = EngineBackend(cfg, device, seed=seed, **engine_kw)
        else:
            from .models.torch_backend import TorchBackend

            be = TorchBackend(cfg, device, dtype=dtype, seed=seed)
        if os.path.isdir(name_or_path):
            sd = checkpoint.load_pretrained_state(name_or_path, cfg)
            if sd is not None:
                be.load_state_dict(sd)
            ts = checkpoint.load_training_state(name_or_path)  # true resume: Adam moments + step counter
            if ts is not None:
                be.load_optimizer_state(ts["optimizer"])
        from .data import try_load_tokenizer

        return SequenceClassifier(cfg, be, tokenizer=try_load_tokenizer(name_or_path))


# ---------------------------------------------------------------------- results files
def write_train_results(path: str, history: History, train_runtime: Optional[float] = None) -> None:
    """``train_results.txt``: one ``key = value`` line per History item (value = list with one
    float per epoch) and, for the distributed script, ``train_runtime = {'train_runtime': s}``
    (scripts/train.py:157-165 — the dict-valued line is the reference's exact format)."""
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "w") as w:
        for k, v in history.history.items():
            w.write("%s = %s\n" % (k, v))
        if train_runtime is not None:
            w.write(f"train_runtime = {{'train_runtime': {round(train_runtime, 4)}}}\n")


def write_eval_results(path: str, result: Dict[str, float]) -> None:
    """``eval_results.txt``: ``key = value`` per metric (scripts/train.py:172-179)."""
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "w") as w:
        for k, v in result.items():
            w.write("%s = %s\n" % (k, v))


def setup_logging() -> logging.Logger:
    """Same handler/format as the reference (scripts/train.py:55-61)."""
    logging.basicConfig(level=logging.getLevelName("INFO"), handlers=[logging.StreamHandler(sys.stdout)],
                        format="%(asctime)s - %(name)s - %(levelname)s - %(message)s", force=True)
    return logging.getLogger("train")
