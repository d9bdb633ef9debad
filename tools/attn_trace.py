"""clock64() timeline of CTA (0,0,0) of the fused attention forward (B=8, H=12, S=512).

Slots (cycles relative to kernel entry of the CTA):
  softmax thread 0:  0 entry | 1 set-up done | 2..5 S chunk c visible | 6 pass 1 done |
                     8..11 pass 2 chunk c handed to the MMA | 12 row sums exchanged | 13 O ready | 14 stored
  control thread:    16 TMA issued | 17 Q,K landed | 18 QK^T issued | 19 V landed | 20..23 P·V chunk c issued
"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from huggingface_sagemaker_tensorflow_distributed_b200.ops import _lib, kernels as K  # noqa: E402

B, H, S = 8, 12, 512
d = H * 64
torch.manual_seed(0)
qkv = torch.randn(B * S, 3 * d, device="cuda").bfloat16()
mask = torch.ones(B, S, dtype=torch.int32, device="cuda")
ctx = torch.empty(B * S, d, dtype=torch.bfloat16, device="cuda")
lse = torch.empty(B, H, S, device="cuda")
st = K.new_step_state("cuda", seed=1)
K.step_prologue(st, 1e-3, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.0)
lib = _lib.load()
lib.hstd_attn_set_trace.argtypes = [C.c_void_p]
lib.hstd_attn_set_trace.restype = None
names = {0: "entry", 1: "setup done", 2: "S0 visible", 3: "S1 visible", 4: "S2 visible", 5: "S3 visible", 6: "pass1 done",
         8: "P0 handed", 9: "P1 handed", 10: "P2 handed", 11: "P3 handed", 12: "sums exchanged", 13: "O ready",
         14: "stored", 16: "ctl: TMA issued", 17: "ctl: Q,K landed", 18: "ctl: QK issued", 19: "ctl: V landed",
         20: "ctl: PV0 issued", 21: "ctl: PV1 issued", 22: "ctl: PV2 issued", 23: "ctl: PV3 issued"}
dO = torch.randn(B * S, d, device="cuda").bfloat16()
dqkv = torch.empty(B * S, 3 * d, dtype=torch.bfloat16, device="cuda")
ws = torch.empty(K.attn_bwd_workspace_bytes(B, H, S), dtype=torch.uint8, device="cuda")
bnames = {24: "entry", 25: "dK/dV staged", 26: "ctl: all stores read"}
for i in range(4):
    bnames.update({32 + 8 * i: f"ctl: Q/dO[{i}] landed", 33 + 8 * i: f"ctl: Pm/dS'[{i}] ready", 34 + 8 * i: f"ctl: MMA2[{i}] complete",
                   36 + 8 * i: f"S/dP[{i}] visible", 37 + 8 * i: f"Pm/dS'[{i}] written", 38 + 8 * i: f"dQ[{i}] visible",
                   39 + 8 * i: f"dQ[{i}] staged"})
for p in (0.1,):
    trace = torch.zeros(64, dtype=torch.int64, device="cuda")
    K.attn_fwd(qkv, ctx, lse, mask, B, H, S, 0.125, p_drop=p, site=1, state=st)
    K.attn_bwd(qkv, ctx, dO, lse, mask, dqkv, ws, B, H, S, 0.125, p_drop=p, site=1, state=st)
    torch.cuda.synchronize()
    lib.hstd_attn_set_trace(trace.data_ptr())
    K.attn_bwd(qkv, ctx, dO, lse, mask, dqkv, ws, B, H, S, 0.125, p_drop=p, site=1, state=st)
    torch.cuda.synchronize()
    lib.hstd_attn_set_trace(None)
    t = trace.cpu().tolist()
    print(f"--- backward p_drop={p} warm L2")
    for k in sorted(bnames, key=lambda k: t[k]):
        if t[k]:
            print(f"  {t[k] - t[24]:7d}  {bnames[k]}")

for p in (0.0, 0.1):
    for warm in (False, True):
        trace = torch.zeros(64, dtype=torch.int64, device="cuda")
        if not warm:
            torch.empty(256 << 20, dtype=torch.uint8, device="cuda").zero_()  # evict L2
        lib.hstd_attn_set_trace(trace.data_ptr())
        K.attn_fwd(qkv, ctx, lse, mask, B, H, S, 0.125, p_drop=p, site=1, state=st)
        torch.cuda.synchronize()
        lib.hstd_attn_set_trace(None)
        t = trace.cpu().tolist()
        print(f"--- p_drop={p} {'warm L2' if warm else 'cold L2'}")
        for k in sorted(names, key=lambda k: t[k]):
            if t[k]:
                print(f"  {t[k] - t[0]:7d}  {names[k]}")
