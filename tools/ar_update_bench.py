"""Isolated timing of the fused all-reduce + AdamW bucket kernel (no concurrent compute): bucket sizes x variants
x block counts, under torchrun with 2/4/8 ranks.  CUDA events, max over ranks, L2 not flushed (state streams
through anyway: 28 B/param)."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import huggingface_sagemaker_tensorflow_distributed_b200.parallel.hvd as hvd  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.ops import kernels as K  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.parallel import comm, p2p  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.parallel.dp_engine import Bucket  # noqa: E402


def timed(fn, iters, dev):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item() * 1e3


def main():
    hvd.init()
    W, R, dev = hvd.size(), hvd.rank(), comm.device()
    total = 32 << 20  # elements (64 MB of bf16)
    grp = p2p.P2PGroup(total, [Bucket(0, 0, total, [])], dev)
    grp.g16.normal_(std=0.01)
    st = K.new_step_state(dev, seed=1)
    K.step_prologue(st, 1e-3, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.0)
    master = torch.randn(total, device=dev)
    m, v = torch.zeros_like(master), torch.zeros_like(master)
    modes = [("two_shot", p2p.AR_TWO_SHOT)] + ([("nvls", p2p.AR_NVLS)] if grp.arena.mc_ptr else [])
    for mb in (2, 8, 16, 32):
        numel = mb * (1 << 20) // 2 // (8 * W) * (8 * W)
        per_rank = numel // W
        line = f"{mb:3d} MB bucket ({per_rank/1e6:.2f} M params/rank):"
        for blocks in (32, 64, 128):
            a = grp.make_args(0, 0, per_rank, p2p.AR_LOCAL, True, st, master, m, v, 0, blocks)
            us = timed(lambda: grp.launch(a), 20, dev)
            line += f" local/b{blocks} {us:6.1f}us"
        for name, mode in modes:
            for upd in (False, True):
                for blocks in (32, 64, 128):
                    a = grp.make_args(0, 0, numel, mode, upd, st, master, m, v, 0, blocks)
                    us = timed(lambda: grp.launch(a), 20, dev)
                    line += f" {name}{'+adamw' if upd else ''}/b{blocks} {us:6.1f}us"
        grp.check_errors()
        if R == 0:
            print(line, flush=True)
    if R == 0:
        n = 8 << 20
        us = timed(lambda: K.adamw_flat(st, grp.g16[:n], master[:n], m[:n], v[:n], grp.w16[:n]), 20, dev)
        print(f"adamw_flat (full grid) {n/1e6:.1f} M params: {us:.1f} us = {n*28/us/1e6:.2f} TB/s")
    else:
        n = 8 << 20
        timed(lambda: K.adamw_flat(st, grp.g16[:n], master[:n], m[:n], v[:n], grp.w16[:n]), 20, dev)
    hvd.shutdown()


main()
