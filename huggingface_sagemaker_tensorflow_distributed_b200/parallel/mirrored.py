"""``tf.distribute.MirroredStrategy``-shaped single-process front-end
(scripts/singe_node_train.py:40-41).

One Python process drives every visible GPU (or the CPU when there is none).  Semantics
kept from the reference: the batch size passed to ``fit`` is the *global* batch and is split
across replicas; there is no LR scaling and no explicit broadcast (replicas are created
identical).  With 0 or 1 device this is plain single-device training — the CPU oracle path
of BASELINE config #1.  With N > 1 GPUs the strategy runs one replica thread per device over
the same engine and the same P2P all-reduce kernels as the multi-process path (peer access
inside one process needs no IPC handles).
"""
from __future__ import annotations

import contextlib
from typing import List

import torch

_CURRENT = None


class MirroredStrategy:
    def __init__(self, devices: List[str] | None = None):
        if devices is None:
            n = torch.cuda.device_count() if torch.cuda.is_available() else 0
            devices = [f"cuda:{i}" for i in range(n)] or ["cpu"]
        self.devices = [torch.device(d) for d in devices]

    @property
    def num_replicas_in_sync(self) -> int:
        return len(self.devices)

    @contextlib.contextmanager
    def scope(self):
        global _CURRENT
        prev, _CURRENT = _CURRENT, self
        try:
            yield self
        finally:
            _CURRENT = prev


def current_strategy() -> "MirroredStrategy | None":
    return _CURRENT
