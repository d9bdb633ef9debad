"""Attention microbenchmark, B=8 H=12 S=512 dh=64 (one DistilBERT / BERT-base layer): the kernels in
the tree against the library baseline F.scaled_dot_product_attention (flash / cuDNN / mem-efficient back
ends, with and without the key-padding mask and dropout), forward and forward+backward, same box, same shape.

CUDA events around one call, L2 flushed before every call, median of ITERS."""
import ctypes as C
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from huggingface_sagemaker_tensorflow_distributed_b200.ops import _lib, kernels as K  # noqa: E402

B, H, S = int(os.environ.get("AB", 8)), int(os.environ.get("AH", 12)), int(os.environ.get("AS", 512))
iters = int(os.environ.get("ITERS", 30))
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
lib = _lib.load()


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


def graph_time(fn, reps=10):
    """Device time per call with no host launch gaps: `reps` calls captured in one CUDA graph (inputs stay
    L2-resident between calls, as they are inside a training step right after the QKV projection)."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3 / reps)
    ts.sort()
    return ts[len(ts) // 2]


fl_fwd = 4.0 * B * H * S * S * 64
stream = torch.cuda.current_stream().cuda_stream
for p in (0.0, 0.1):
    f = timeit(lambda: K.attn_fwd(qkv, ctx, lse, mask, B, H, S, 0.125, p_drop=p, site=1, state=st))
    b = timeit(lambda: K.attn_bwd(qkv, ctx, dO, lse, mask, dqkv, ws, B, H, S, 0.125, p_drop=p, site=1, state=st))
    print(f"ours p_drop={p}: fwd {f:.1f} us ({fl_fwd/f/1e6:.0f} TFLOP/s)  bwd(+delta+reduce) {b:.1f} us "
          f"({2.5*fl_fwd/b/1e6:.0f} TFLOP/s)")

    def ours_fb():
        # launched on the capturing stream
        K.attn_fwd(qkv, ctx, lse, mask, B, H, S, 0.125, p_drop=p, site=1, state=st)
        K.attn_bwd(qkv, ctx, dO, lse, mask, dqkv, ws, B, H, S, 0.125, p_drop=p, site=1, state=st)

    try:
        gf = graph_time(lambda: K.attn_fwd(qkv, ctx, lse, mask, B, H, S, 0.125, p_drop=p, site=1, state=st))
        gfb = graph_time(ours_fb)
        print(f"ours p_drop={p} (graph x10, L2-warm): fwd {gf:.1f} us  fwd+bwd {gfb:.1f} us")
    except Exception as e:  # noqa: BLE001
        print(f"ours graph timing failed: {type(e).__name__}: {str(e)[:200]}")

# ---- library baseline: torch SDPA (flash / cuDNN / mem-efficient back ends as torch selects), [B,H,S,64] bf16
q, k, v = (torch.randn(B, H, S, 64, device="cuda", dtype=torch.bfloat16, requires_grad=True) for _ in range(3))
go = torch.randn(B, H, S, 64, device="cuda", dtype=torch.bfloat16)
bias = torch.zeros(B, 1, 1, S, device="cuda", dtype=torch.bfloat16)
from torch.nn.attention import SDPBackend, sdpa_kernel  # noqa: E402

for name, backends, am in (("flash (no mask)", [SDPBackend.FLASH_ATTENTION], None),
                           ("cudnn (no mask)", [SDPBackend.CUDNN_ATTENTION], None),
                           ("cudnn (key-padding bias)", [SDPBackend.CUDNN_ATTENTION], bias),
                           ("mem-efficient (key-padding bias)", [SDPBackend.EFFICIENT_ATTENTION], bias),
                           ("default (key-padding bias)", None, bias)):
    for p in (0.0, 0.1):
        try:
            def fwd():
                if backends is None:
                    return F.scaled_dot_product_attention(q, k, v, attn_mask=am, dropout_p=p)
                with sdpa_kernel(backends):
                    return F.scaled_dot_product_attention(q, k, v, attn_mask=am, dropout_p=p)

            def fwdbwd():
                torch.autograd.grad(fwd(), (q, k, v), go)

            with torch.no_grad():
                tf = timeit(fwd)
            tfb = timeit(fwdbwd)
            print(f"SDPA {name} p_drop={p}: fwd {tf:.1f} us  fwd+bwd {tfb:.1f} us  (bwd ~{tfb - tf:.1f} us)  [single eager calls, L2 flushed]")
            try:
                with torch.no_grad():
                    gf = graph_time(fwd)
                gfb = graph_time(fwdbwd)
                print(f"SDPA {name} p_drop={p} (graph x10, L2-warm): fwd {gf:.1f} us  fwd+bwd {gfb:.1f} us")
            except Exception as e:  # noqa: BLE001
                print(f"SDPA {name} p_drop={p}: graph timing unavailable ({type(e).__name__}: {str(e)[:100]})")
                torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001
            print(f"SDPA {name} p_drop={p}: unavailable ({type(e).__name__}: {str(e)[:100]})")
