"""Gradient all-reduce bandwidth sweep, 1 KB – 1 GB bf16, our NVLink kernels vs NCCL
(BASELINE config 5).  Run under torchrun with 2/4/8 ranks; rank 0 writes
gpurun_out/allreduce_sweep_<N>gpu.{txt,json}.  Times are CUDA events on the launching
stream, max over ranks; busbw = algbw * 2(W-1)/W."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import huggingface_sagemaker_tensorflow_distributed_b200.parallel.hvd as hvd  # noqa: E402
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
    return t.item() * 1e3  # us


def check_values(grp, numel, mode, name, W, R, dev):
    """One averaging all-reduce of rank-dependent data against the gathered fp32 mean, and bit-equality of the
    result on every rank (the timing loops below say nothing about correctness)."""
    g = torch.Generator(device=dev).manual_seed(1234 + R)
    grp.g16[:numel].copy_((torch.randn(numel, device=dev, generator=g) * (1 + R)).bfloat16())
    grp.w16[:numel].zero_()
    torch.cuda.synchronize()
    dist.barrier()
    parts = [torch.empty(numel, dtype=torch.bfloat16, device=dev) for _ in range(W)]
    dist.all_gather(parts, grp.g16[:numel].contiguous())
    want = torch.stack([p.float() for p in parts]).sum(0) / W
    del parts
    dist.barrier()
    grp.launch(grp.make_args(0, 0, numel, mode, update=False, blocks=64))
    torch.cuda.synchronize()
    dist.barrier()
    got = (grp.w16 if mode == p2p.AR_ONE_SHOT else grp.g16)[:numel]
    err = (got.float() - want).abs().max().item()
    assert err <= 2 ** -7 * want.abs().max().item() + 1e-6, (name, numel, err)
    outs = [torch.empty_like(got) for _ in range(W)]
    dist.all_gather(outs, got.contiguous())
    assert all(torch.equal(outs[0], o) for o in outs), f"{name}: ranks hold different results at {numel} elements"
    grp.check_errors()
    grp.g16[:numel].normal_()


def main():
    hvd.init()
    W, R, dev = hvd.size(), hvd.rank(), comm.device()
    max_bytes = int(os.environ.get("SWEEP_MAX_BYTES", str(1 << 30)))
    sizes = [1 << k for k in range(10, 31) if (1 << k) <= max_bytes]
    check_max_bytes = int(os.environ.get("SWEEP_CHECK_MAX_BYTES", str(16 << 20)))
    total = max_bytes // 2
    grp = p2p.P2PGroup(total, [Bucket(0, 0, total, [])], dev)
    grp.g16.normal_()
    rows = []
    modes = [("two_shot", p2p.AR_TWO_SHOT), ("one_shot", p2p.AR_ONE_SHOT)]
    if grp.arena.mc_ptr:
        modes.append(("nvls", p2p.AR_NVLS))
    nccl_buf = torch.randn(total, device=dev).bfloat16()
    for nbytes in sizes:
        numel = max(nbytes // 2, 8 * W) // (8 * W) * (8 * W)
        iters = 50 if nbytes <= (16 << 20) else (20 if nbytes <= (256 << 20) else 8)
        row = {"bytes": numel * 2}
        t = nccl_buf[:numel]
        row["nccl_us"] = timed(lambda: dist.all_reduce(t), iters, dev)
        for name, mode in modes:
            if mode == p2p.AR_ONE_SHOT and nbytes > (64 << 20):
                continue
            if nbytes <= check_max_bytes:
                check_values(grp, numel, mode, name, W, R, dev)
                row["values_checked"] = True
            best = None
            for blocks, tail in ((8, False), (32, False), (128, False), (32, True), (128, True)):
                if blocks * 128 * (4 if tail else 1) * 8 > numel * 4 and blocks > 8:
                    continue
                args = grp.make_args(0, 0, numel, mode, update=False, blocks=blocks, tail=tail)
                us = timed(lambda: grp.launch(args), iters, dev)
                if best is None or us < best[0]:
                    best = (us, f"{blocks}x{512 if tail else 128}")
            row[name + "_us"], row[name + "_blocks"] = best
        grp.check_errors()
        rows.append(row)
        if R == 0:
            f = 2 * (W - 1) / W
            msg = f"{row['bytes']:>11d} B | nccl {row['nccl_us']:9.1f} us {row['bytes']*f/row['nccl_us']/1e3:7.1f} GB/s"
            for name, _ in modes:
                if name + "_us" in row:
                    us = row[name + "_us"]
                    msg += f" | {name} {us:9.1f} us {row['bytes']*f/us/1e3:7.1f} GB/s ({row[name+'_blocks']})"
            print(msg, flush=True)
    if R == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/allreduce_sweep_{W}gpu.json", "w") as fjs:
            json.dump({"world": W, "arena": grp.arena.kind, "nvls": bool(grp.arena.mc_ptr), "rows": rows}, fjs, indent=1)
    hvd.shutdown()


if __name__ == "__main__":
    main()
