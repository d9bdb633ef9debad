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


def main():
    hvd.init()
    W, R, dev = hvd.size(), hvd.rank(), comm.device()
    max_bytes = int(os.environ.get("SWEEP_MAX_BYTES", str(1 << 30)))
    sizes = [1 << k for k in range(10, 31) if (1 << k) <= max_bytes]
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
            best = None
            for blocks in (8, 32, 64, 128):
                if blocks * 512 * 8 > numel * 4 and blocks > 8:
                    continue
                args = grp.make_args(0, 0, numel, mode, update=False, blocks=blocks)
                us = timed(lambda: grp.launch(args), iters, dev)
                if best is None or us < best[0]:
                    best = (us, blocks)
            row[name + "_us"], row[name + "_blocks"] = best
        grp.check_errors()
        rows.append(row)
        if R == 0:
            f = 2 * (W - 1) / W
            msg = f"{row['bytes']:>11d} B | nccl {row['nccl_us']:9.1f} us {row['bytes']*f/row['nccl_us']/1e3:7.1f} GB/s"
            for name, _ in modes:
                if name + "_us" in row:
                    us = row[name + "_us"]
                    msg += f" | {name} {us:9.1f} us {row['bytes']*f/us/1e3:7.1f} GB/s (b{row[name+'_blocks']})"
            print(msg, flush=True)
    if R == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/allreduce_sweep_{W}gpu.json", "w") as fjs:
            json.dump({"world": W, "arena": grp.arena.kind, "nvls": bool(grp.arena.mc_ptr), "rows": rows}, fjs, indent=1)
    hvd.shutdown()


if __name__ == "__main__":
    main()
