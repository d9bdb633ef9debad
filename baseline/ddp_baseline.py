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


def build(cfg, device, world: int, lr: float, sdpa: bool = True):
    from huggingface_sagemaker_tensorflow_distributed_b200.models.reference import EncoderForSequenceClassification

    torch.manual_seed(42)
    model = EncoderForSequenceClassification(cfg, sdpa=sdpa).to(device)
    if world > 1:
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[device.index], bucket_cap_mb=25,
                                                          gradient_as_bucket_view=True)
    opt = torch.optim.AdamW(model.parameters(), lr=lr, eps=1e-7, weight_decay=0.0, fused=True)
    return model, opt


def train_step(model, opt, ids, mask, labels):
    opt.zero_grad(set_to_none=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        logits = model(ids, mask)
    loss = F.cross_entropy(logits.float(), labels)
    loss.backward()
    opt.step()
    return loss
