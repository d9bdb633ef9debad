"""Keras-semantics Adam / AdamW and LR schedules.

The reference builds ``tf.keras.optimizers.Adam(learning_rate=lr * hvd.size())``
(scripts/train.py:112-113; no scaling in scripts/singe_node_train.py:78).  Keras Adam is not
``torch.optim.Adam``: ε defaults to 1e-7 and sits outside the bias correction,

    lr_t = lr · sqrt(1 − β2^t) / (1 − β1^t)
    w   −= lr_t · m / (sqrt(v) + ε)

``weight_decay`` (decoupled, ``w −= lr · wd · w``) and the linear-decay schedule are the
north-star additions; ``weight_decay=0`` + ``constant`` reproduces the reference exactly.
The same formulas are implemented on the device in csrc/optimizer.cu and fused into the
gradient all-reduce (csrc/allreduce_sm100.cu); this module is the fp32 oracle / CPU path.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Iterable, Optional

import torch


@dataclass
class LRSchedule:
    """constant | linear (linear warm-up to ``base_lr`` then linear decay to 0 at total_steps)."""
    base_lr: float
    kind: str = "constant"
    total_steps: int = 0
    warmup_steps: int = 0

    def __call__(self, step: int) -> float:
        """Learning rate for optimizer step ``step`` (0-based)."""
        if self.kind == "constant":
            return self.base_lr
        if self.kind != "linear":
            raise ValueError(f"unknown lr schedule {self.kind!r}")
        if self.warmup_steps > 0 and step < self.warmup_steps:
            return self.base_lr * float(step + 1) / float(self.warmup_steps)
        total = max(self.total_steps, self.warmup_steps + 1)
        return self.base_lr * max(0.0, float(total - step) / float(total - self.warmup_steps))


def scale_learning_rate(lr: float, world_size: int) -> float:
    """Linear LR scaling with the number of workers (scripts/train.py:112)."""
    return lr * world_size


class KerasAdamW(torch.optim.Optimizer):
    def __init__(self, params: Iterable, learning_rate: float = 1e-3, beta_1: float = 0.9, beta_2: float = 0.999,
                 epsilon: float = 1e-7, weight_decay: float = 0.0, schedule: Optional[LRSchedule] = None):
        defaults = dict(lr=learning_rate, beta_1=beta_1, beta_2=beta_2, epsilon=epsilon, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.schedule = schedule or LRSchedule(learning_rate)
        self.iterations = 0  # Keras name for the step counter
        self.distributed = False  # set by parallel.hvd.DistributedOptimizer

    @property
    def learning_rate(self) -> float:
        return self.schedule(self.iterations)

    def hyper(self) -> dict:
        g = self.param_groups[0]
        return dict(beta_1=g["beta_1"], beta_2=g["beta_2"], epsilon=g["epsilon"], weight_decay=g["weight_decay"],
                    schedule=self.schedule)

    @torch.no_grad()
    def step(self, closure=None):
        t = self.iterations + 1
        lr = self.schedule(self.iterations)
        for g in self.param_groups:
            b1, b2, eps, wd = g["beta_1"], g["beta_2"], g["epsilon"], g["weight_decay"]
            lr_t = lr * math.sqrt(1.0 - b2 ** t) / (1.0 - b1 ** t)
            for p in g["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if not st:
                    st["m"] = torch.zeros_like(p, dtype=torch.float32)
                    st["v"] = torch.zeros_like(p, dtype=torch.float32)
                grad = p.grad.float()
                st["m"].mul_(b1).add_(grad, alpha=1.0 - b1)
                st["v"].mul_(b2).addcmul_(grad, grad, value=1.0 - b2)
                if wd != 0.0:
                    p.mul_(1.0 - lr * wd)
                p.addcdiv_(st["m"], st["v"].sqrt().add_(eps), value=-lr_t)
        self.iterations = t
        return None


class Adam:
    """``tf.keras.optimizers.Adam``-shaped optimizer *description*: like Keras, it is built
    before the model and without parameters (scripts/train.py:113); ``model.compile`` binds
    it — to ``KerasAdamW`` on the PyTorch path, or to the fused device optimizer inside the
    all-reduce kernel on the B200 path."""

    def __init__(self, learning_rate: float = 1e-3, beta_1: float = 0.9, beta_2: float = 0.999,
                 epsilon: float = 1e-7, weight_decay: float = 0.0, schedule: Optional[LRSchedule] = None,
                 lr_schedule: str = "constant", warmup_steps: int = 0):
        self.learning_rate = float(learning_rate)
        self.beta_1, self.beta_2, self.epsilon = float(beta_1), float(beta_2), float(epsilon)
        self.weight_decay = float(weight_decay)
        self.schedule = schedule  # resolved at fit() time when total steps are known
        self.lr_schedule, self.warmup_steps = lr_schedule, int(warmup_steps)
        self.distributed = False  # set by parallel.hvd.DistributedOptimizer
        self.iterations = 0

    def resolve_schedule(self, total_steps: int) -> LRSchedule:
        if self.schedule is None:
            self.schedule = LRSchedule(self.learning_rate, self.lr_schedule, total_steps, self.warmup_steps)
        return self.schedule

    def bind(self, params: Iterable, total_steps: int = 0) -> KerasAdamW:
        opt = KerasAdamW(params, self.learning_rate, self.beta_1, self.beta_2, self.epsilon, self.weight_decay,
                         self.resolve_schedule(total_steps))
        opt.distributed = self.distributed
        opt.iterations = self.iterations
        return opt


AdamW = Adam
