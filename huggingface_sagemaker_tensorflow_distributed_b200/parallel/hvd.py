"""Horovod-Keras-shaped distribution API (SURVEY.md §1 layer D, §2.1 C8, C9, C13, C14).

The reference does ``import horovod.tensorflow.keras as hvd`` — or the API-compatible
``smdistributed.dataparallel.tensorflow.keras`` when ``SM_FRAMEWORK_PARAMS`` enables it
(scripts/train.py:16-21) — and then uses exactly: ``init``, ``local_rank``, ``rank``,
``size``, ``DistributedOptimizer`` and ``callbacks.BroadcastGlobalVariablesCallback``
(scripts/train.py:24,31,112-114,127-134,152).  This module offers the same names:

    import huggingface_sagemaker_tensorflow_distributed_b200.parallel.hvd as hvd
    hvd.init()
    opt = hvd.DistributedOptimizer(Adam(model.parameters(), lr * hvd.size()))
    model.fit(ds, callbacks=[hvd.callbacks.BroadcastGlobalVariablesCallback(0)])

``DistributedOptimizer`` marks the optimizer so the training runtime averages gradients
across ranks before the update: on B200 that is the fused P2P reduce+AdamW kernel, on CPU a
gloo all-reduce.  ``backend_name()`` reports which front-end ("smddp" or "horovod") the
environment selected; both drive the same engine.
"""
from __future__ import annotations

from types import SimpleNamespace

from ..config import is_sagemaker_dp_enabled
from . import comm


def init() -> None:
    comm.init()


def shutdown() -> None:
    comm.shutdown()


def rank() -> int:
    return comm.rank()


def size() -> int:
    return comm.size()


def local_rank() -> int:
    return comm.local_rank()


def local_size() -> int:
    return comm.state().local_size


def is_initialized() -> bool:
    return comm.state().initialized


def backend_name() -> str:
    return "smddp" if is_sagemaker_dp_enabled() else "horovod"


def DistributedOptimizer(optimizer, **_ignored):
    """Marks ``optimizer`` as data-parallel: gradients are all-reduce-averaged across ranks
    before ``apply`` (scripts/train.py:114).  Returns the same object."""
    optimizer.distributed = True
    return optimizer


class BroadcastGlobalVariablesCallback:
    """Broadcast model + optimizer state from ``root_rank`` at the start of ``fit``
    (scripts/train.py:127-134)."""

    def __init__(self, root_rank: int = 0):
        self.root_rank = root_rank

    def on_train_begin(self, model) -> None:
        model.broadcast_state(self.root_rank)


class MetricAverageCallback:
    """Average epoch metrics over ranks (the reference omits it, so it is opt-in)."""

    def on_epoch_end(self, model, logs: dict) -> None:
        import torch

        if comm.size() > 1:
            keys = sorted(logs)
            t = torch.tensor([float(logs[k]) for k in keys], dtype=torch.float64, device=comm.device())
            comm.allreduce_mean_(t)
            for k, v in zip(keys, t.tolist()):
                logs[k] = v


callbacks = SimpleNamespace(BroadcastGlobalVariablesCallback=BroadcastGlobalVariablesCallback,
                            MetricAverageCallback=MetricAverageCallback)
