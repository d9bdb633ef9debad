"""Microbenchmark of the fused attention kernels (B=8, H=12, S=512, dh=64 = DistilBERT/BERT-base layer)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from huggingface_sagemaker_tensorflow_distributed_b200.ops import kernels as K  # noqa: E402

B, H, S = int(os.environ.get("AB", 8)), int(os.environ.get("AH", 12)), int(os.environ.get("AS", 512))
iters = int(os.environ.get("ITERS", 20))
d = H * 64
torch.manual_seed(0)
qkv = torch.randn(B * S, 3 * d, device="cuda").bfloat16()
mask = torch.ones(B, S, dtype=torch.int32, device="cuda")
dO = torch.randn(B * S, d, device="cuda").bfloat16()
ctx = torch.empty(B * S, d, dtype=torch.bfloat16, device="cuda")
lse = torch.empty(B, H, S, device="cuda")
dqkv = torch.empty(B * S, 3 * d, dtype=torch.bfloat16, device="cuda")
ws = torch.empty(K.attn_bwd_workspace_bytes(B, H, S), dtype=torch.uint8, device="cuda")
st = K.new_step_state("cuda", seed=1)
K.step_prologue(st, 1e-3, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()  # evict L2
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


fl_fwd = 4.0 * B * H * S * S * 64
for p in (0.0, 0.1):
    f = timeit(lambda: K.attn_fwd(qkv, ctx, lse, mask, B, H, S, 0.125, p_drop=p, site=1, state=st))
    b = timeit(lambda: K.attn_bwd(qkv, ctx, dO, lse, mask, dqkv, ws, B, H, S, 0.125, p_drop=p, site=1, state=st))
    print(f"p_drop={p}: fwd {f:.1f} us ({fl_fwd/f/1e6:.0f} TFLOP/s)  bwd(+delta+reduce) {b:.1f} us ({2.5*fl_fwd/b/1e6:.0f} TFLOP/s)")
