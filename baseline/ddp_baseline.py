"""Same-workload library baseline: PyTorch DDP + NCCL all-reduce + cuBLAS GEMMs + fused SDPA
attention + torch's fused AdamW, bf16 autocast with fp32 master weights.  This is the
"reference NCCL baseline" BASELINE.md falls back to because the TF/Horovod reference cannot
be installed offline.  None of this framework's kernels run here.

Used through ``python bench.py --impl ddp`` (same timing contract as the product arm).
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn.functional as F


def build(cfg, device, world: int, lr: float, sdpa: bool = True, capturable: bool = False):
    from huggingface_sagemaker_tensorflow_distributed_b200.models.reference import EncoderForSequenceClassification

    torch.manual_seed(42)
    model = EncoderForSequenceClassification(cfg, sdpa=sdpa).to(device)
    if world > 1 and capturable:
        # CUDA-graph arm: DistributedDataParallel's reducer does not survive stream capture here
        # (cudaErrorStreamCaptureImplicit), plain NCCL collectives do.  Gradients are views into ONE flat buffer
        # that is all-reduced with NCCL after backward (no overlap with backward: noted in the bench line).
        params = [p for p in model.parameters()]
        flat = torch.zeros(sum(p.numel() for p in params), device=device)
        off = 0
        for p in params:
            p.grad = flat[off:off + p.numel()].view_as(p)
            off += p.numel()
        model._flat_grad, model._world = flat, world
        for p in params:  # rank 0's initial weights everywhere, as DDP's constructor does
            dist.broadcast(p.data, src=0)
    elif world > 1:
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[device.index], bucket_cap_mb=25,
                                                          gradient_as_bucket_view=True)
    opt = torch.optim.AdamW(model.parameters(), lr=lr, eps=1e-7, weight_decay=0.0, fused=True, capturable=capturable)
    return model, opt


def train_step(model, opt, ids, mask, labels):
    flat = getattr(model, "_flat_grad", None)
    if flat is None:
        opt.zero_grad(set_to_none=True)
    else:
        flat.zero_()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        logits = model(ids, mask)
    loss = F.cross_entropy(logits.float(), labels)
    loss.backward()
    if flat is not None:
        dist.all_reduce(flat)
        flat.mul_(1.0 / model._world)
    opt.step()
    return loss


class GraphedStep:
    """The same training step captured ONCE into a CUDA graph (forward, backward, DDP's NCCL all-reduces, fused
    AdamW) and replayed: the library baseline without Python / launch latency, which is what a compiled TF graph
    (``model.fit``, reference scripts/train.py:145) is the counterpart of.  Inputs are copied into static
    tensors before each replay.  DDP asks for >= 11 eager iterations on a side stream before capture."""

    def __init__(self, model, opt, ids, mask, labels, warmup: int = 12):
        self.model, self.opt = model, opt
        self.ids, self.mask, self.labels = ids.clone(), mask.clone(), labels.clone()
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(warmup):
                train_step(model, opt, self.ids, self.mask, self.labels)
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        if getattr(model, "_flat_grad", None) is None:
            opt.zero_grad(set_to_none=True)
        # thread_local: the NCCL watchdog thread polls events while we capture; the default (global) mode
        # treats that as an illegal call during capture (cudaErrorStreamCaptureImplicit)
        with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):
            self.loss = train_step(model, opt, self.ids, self.mask, self.labels)
        torch.cuda.synchronize()

    def __call__(self, ids, mask, labels):
        self.ids.copy_(ids, non_blocking=True)
        self.mask.copy_(mask, non_blocking=True)
        self.labels.copy_(labels, non_blocking=True)
        self.graph.replay()
        return self.loss
