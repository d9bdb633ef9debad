"""GEMM correctness + speed sweep (run on the B200 box).  Writes gpurun_out/gemm_check.txt."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from huggingface_sagemaker_tensorflow_distributed_b200.ops import gemm as G  # noqa: E402

os.makedirs("gpurun_out", exist_ok=True)
out = open("gpurun_out/gemm_check.txt", "w")


def log(*a):
    s = " ".join(str(x) for x in a)
    print(s, flush=True)
    out.write(s + "\n")
    out.flush()


def ref(A, B, a_t, b_t):
    Af = A.float().transpose(-1, -2) if a_t else A.float()
    Bf = B.float() if b_t else B.float().transpose(-1, -2)
    return Af @ Bf


def run_case(M, N, K, a_t, b_t, bn, nb=0, splits=1):
    torch.manual_seed(M + N + K)
    sh = (nb,) if nb else ()
    A = (torch.randn(*sh, *((K, M) if a_t else (M, K)), device="cuda") * 0.5).bfloat16()
    B = (torch.randn(*sh, *((K, N) if b_t else (N, K)), device="cuda") * 0.5).bfloat16()
    try:
        Cc = G.gemm(A, B, a_t=a_t, b_t=b_t, block_n=bn, splits=splits)
        torch.cuda.synchronize()
    except Exception as e:  # noqa: BLE001
        log(f"FAIL M={M} N={N} K={K} a_t={a_t} b_t={b_t} bn={bn} nb={nb} splits={splits}: {e}")
        return False
    R = ref(A, B, a_t, b_t)
    err = (Cc.float() - R).abs().max().item()
    scale = R.abs().max().item()
    ok = err <= 0.02 * scale + 1e-2
    log(f"{'ok ' if ok else 'BAD'} M={M} N={N} K={K} a_t={int(a_t)} b_t={int(b_t)} bn={bn} nb={nb} splits={splits} "
        f"maxerr={err:.4f} refmax={scale:.2f}")
    return ok


def _graph_time(fn, iters):
    """GPU time per call with host launch overhead removed: `iters` calls captured in one CUDA graph."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def bench(M, N, K, a_t, b_t, bn, nb=0, iters=20, **kw):
    sh = (nb,) if nb else ()
    A = torch.randn(*sh, *((K, M) if a_t else (M, K)), device="cuda").bfloat16()
    B = torch.randn(*sh, *((K, N) if b_t else (N, K)), device="cuda").bfloat16()
    Cc = G.gemm(A, B, a_t=a_t, b_t=b_t, block_n=bn, **kw)
    ms = _graph_time(lambda: G.gemm(A, B, a_t=a_t, b_t=b_t, block_n=bn, out=Cc, **kw), iters)
    fl = 2.0 * M * N * K * max(nb, 1)
    Af = A.transpose(-1, -2) if a_t else A
    Bf = B if b_t else B.transpose(-1, -2)
    Cr = torch.matmul(Af, Bf)
    ms_ref = _graph_time(lambda: torch.matmul(Af, Bf, out=Cr), iters)
    log(f"bench M={M} N={N} K={K} a_t={int(a_t)} b_t={int(b_t)} bn={bn} nb={nb} {kw}: {ms*1e3:.1f} us "
        f"{fl/ms/1e9:.0f} TFLOP/s | cuBLAS {ms_ref*1e3:.1f} us {fl/ms_ref/1e9:.0f} TFLOP/s")


def main():
    t0 = time.time()
    allok = True
    # smallest first: a single tile, single k-block
    for (a_t, b_t) in [(False, False), (False, True), (True, False), (True, True)]:
        for bn in (64, 128, 256):
            allok &= run_case(128, bn, 64, a_t, b_t, bn)
    for (a_t, b_t) in [(False, False), (False, True), (True, False), (True, True)]:
        for bn in (64, 128, 256):
            allok &= run_case(256, 512, 256, a_t, b_t, bn)
            allok &= run_case(384, 768, 192, a_t, b_t, bn)
    # edges: M not multiple of 128, N not multiple of BN, K not multiple of 64
    allok &= run_case(200, 328, 136, False, False, 128)
    allok &= run_case(200, 328, 136, False, True, 128)
    allok &= run_case(200, 328, 136, True, True, 128)
    # many tiles per CTA (persistence + TMEM double buffering)
    allok &= run_case(4096, 2304, 768, False, False, 256)
    allok &= run_case(4096, 768, 3072, False, True, 128)
    allok &= run_case(768, 3072, 4096, True, True, 128)
    # batched (attention shapes)
    allok &= run_case(512, 512, 64, False, False, 128, nb=96)
    allok &= run_case(512, 64, 512, False, True, 64, nb=96)
    allok &= run_case(512, 64, 512, True, True, 64, nb=96)
    # split-K
    allok &= run_case(768, 768, 4096, True, True, 128, splits=4)
    allok &= run_case(2304, 768, 4096, True, True, 128, splits=2)
    log("ALL OK" if allok else "SOME FAILED", f"({time.time()-t0:.1f}s)")
    if not allok and os.environ.get("GEMM_BENCH_ANYWAY") != "1":
        return 1
    for cm in (1, 0):  # 1 = single-CTA tiles, 0 = auto (2-SM pairs)
        log(f"--- cluster_m={cm}")
        for bn in (128, 256):
            bench(4096, 2304, 768, False, False, bn, cluster_m=cm)   # QKV fwd
            bench(4096, 3072, 768, False, False, bn, cluster_m=cm)   # FFN1 fwd
            bench(4096, 768, 3072, False, False, bn, cluster_m=cm)   # FFN2 fwd
            bench(4096, 768, 768, False, False, bn, cluster_m=cm)    # out-proj fwd
        bench(8192, 8192, 8192, False, False, 256, iters=5, cluster_m=cm)
        bench(4096, 768, 3072, False, True, 256, cluster_m=cm)       # FFN1 dgrad
        bench(4096, 3072, 768, False, True, 256, cluster_m=cm)       # FFN2 dgrad
        bench(3072, 768, 4096, True, True, 128, cluster_m=cm)        # FFN1 wgrad
        bench(3072, 768, 4096, True, True, 256, cluster_m=cm)
        bench(768, 3072, 4096, True, True, 256, cluster_m=cm)        # FFN2 wgrad
        bench(2304, 768, 4096, True, True, 256, cluster_m=cm)        # QKV wgrad
        bench(768, 768, 4096, True, True, 128, splits=4, cluster_m=cm)
        bench(768, 768, 4096, True, True, 256, splits=4, cluster_m=cm)
    return 0


if __name__ == "__main__":
    sys.exit(main())
