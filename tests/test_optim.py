"""Keras-Adam semantics: eps=1e-7 outside the bias correction."""
import math

import pytest
import torch

from huggingface_sagemaker_tensorflow_distributed_b200.optim import Adam, KerasAdamW, LRSchedule, scale_learning_rate


def _keras_adam_numpy(w, grads, lr, b1=0.9, b2=0.999, eps=1e-7, wd=0.0):
    m = v = 0.0
    for t, g in enumerate(grads, 1):
        m = b1 * m + (1 - b1) * g
        v = b2 * v + (1 - b2) * g * g
        lr_t = lr * math.sqrt(1 - b2 ** t) / (1 - b1 ** t)
        w = w * (1 - lr * wd) - lr_t * m / (math.sqrt(v) + eps)
    return w


@pytest.mark.parametrize("wd", [0.0, 0.01])
def test_matches_keras_formula(wd):
    p = torch.nn.Parameter(torch.tensor([0.5], dtype=torch.float64))
    opt = KerasAdamW([p], learning_rate=1e-2, weight_decay=wd)
    grads = [0.3, -0.2, 1e-9, 0.7]
    for g in grads:
        p.grad = torch.tensor([g], dtype=torch.float64)
        opt.step()
    assert p.item() == pytest.approx(_keras_adam_numpy(0.5, grads, 1e-2, wd=wd), rel=1e-6)


def test_differs_from_torch_adam_for_tiny_grads():
    # with |g| ~ eps the epsilon placement matters: this is what makes it "Keras" Adam
    p1 = torch.nn.Parameter(torch.tensor([1.0]))
    p2 = torch.nn.Parameter(torch.tensor([1.0]))
    o1, o2 = KerasAdamW([p1], learning_rate=1e-3), torch.optim.Adam([p2], lr=1e-3, eps=1e-7)
    for _ in range(3):
        p1.grad = torch.tensor([1e-7]); p2.grad = torch.tensor([1e-7])
        o1.step(); o2.step()
    assert abs(p1.item() - p2.item()) > 1e-6


def test_linear_schedule_and_scaling():
    s = LRSchedule(1.0, "linear", total_steps=10, warmup_steps=2)
    assert s(0) == pytest.approx(0.5) and s(1) == pytest.approx(1.0)
    assert s(2) == pytest.approx(1.0) and s(6) == pytest.approx(0.5) and s(10) == 0.0
    assert LRSchedule(3e-5)(123) == 3e-5
    assert scale_learning_rate(5e-5, 8) == pytest.approx(4e-4)  # scripts/train.py:112


def test_spec_binds_and_keeps_distributed_flag():
    spec = Adam(learning_rate=1e-3, lr_schedule="linear")
    spec.distributed = True
    p = torch.nn.Parameter(torch.zeros(2))
    opt = spec.bind([p], total_steps=100)
    assert opt.distributed and opt.schedule.total_steps == 100 and opt.defaults["epsilon"] == 1e-7
