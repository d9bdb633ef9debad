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
    if world > 1:
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[device.index], bucket_cap_mb=25,
                                                          gradient_as_bucket_view=True)
    opt = torch.optim.AdamW(model.parameters(), lr=lr, eps=1e-7, weight_decay=0.0, fused=True, capturable=capturable)
    return model, opt


def train_step(model, opt, ids, mask, labels):
    opt.zero_grad(set_to_none=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        logits = model(ids, mask)
    loss = F.cross_entropy(logits.float(), labels)
    loss.backward()
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
        opt.zero_grad(set_to_none=True)
        with torch.cuda.graph(self.graph):
            self.loss = train_step(model, opt, self.ids, self.mask, self.labels)
        torch.cuda.synchronize()

    def __call__(self, ids, mask, labels):
        self.ids.copy_(ids, non_blocking=True)
        self.mask.copy_(mask, non_blocking=True)
        self.labels.copy_(labels, non_blocking=True)
        self.graph.replay()
        return self.loss
