"""Isolated timing of the memory-bound row kernels at the model's shapes (M = 4096 tokens): LayerNorm forward /
backward (dropout on), bias-gradient column sums.  Each kernel 20x inside one CUDA graph (no launch gaps; data
stays L2-resident as it is inside a training step), median of 10 replays; HBM-bound time for reference."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from huggingface_sagemaker_tensorflow_distributed_b200.ops import kernels as K  # noqa: E402

M = int(os.environ.get("M", 4096))
dev = "cuda"
st = K.new_step_state(dev, seed=1)
K.step_prologue(st, 1e-3, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.0)


def graph_us(fn, reps=20):
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
    ts = []
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3 / reps)
    return sorted(ts)[len(ts) // 2]


for d in (768, 1024):
    bf = lambda *s: torch.randn(*s, device=dev).bfloat16()  # noqa: E731
    a, resid, pre, y, dy, dpre, da = (bf(M, d) for _ in range(7))
    gamma, beta, dg, db, dbias = (bf(d) for _ in range(5))
    mean, rstd = torch.zeros(M, device=dev), torch.ones(M, device=dev)
    acc = torch.zeros(K.ln_bwd_partials_floats(d), device=dev)
    for p in (0.0, 0.1):
        t = graph_us(lambda: K.ln_fwd(a=a, resid=resid, pre=pre, y=y, mean=mean, rstd=rstd, gamma=gamma, beta=beta,
                                      eps=1e-12, state=st, p_in=p, site_in=3))
        print(f"d={d} p={p}: ln_fwd (dropout+residual+LN) {t:6.2f} us   [{4*M*d*2/6.4e6:.1f} us at 6.4 TB/s for 4 x M x d bf16]")
        t = graph_us(lambda: K.ln_bwd(dy=dy, pre=pre, mean=mean, rstd=rstd, gamma=gamma, dpre=dpre, da=da if p > 0 else None,
                                      dgamma=dg, dbeta=db, dbias=dbias, partials=acc, state=st, p_in=p, site_in=3))
        n = 4 + (p > 0)
        print(f"d={d} p={p}: ln_bwd (rows + column sums)      {t:6.2f} us   [{n*M*d*2/6.4e6:.1f} us at 6.4 TB/s]")
for N in (768, 2304, 3072):
    x = torch.randn(M, N, device=dev).bfloat16()
    out = torch.empty(N, dtype=torch.bfloat16, device=dev)
    part = torch.empty(K.colsum_partials_floats(N), device=dev)
    t = graph_us(lambda: K.colsum(x, out, part))
    print(f"colsum N={N}: {t:6.2f} us   [{M*N*2/6.4e6:.1f} us at 6.4 TB/s]")
