"""Per-k-block clock64() timeline of CTA 0 of one GEMM: who waits on whom in the mainloop."""
import ctypes, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from huggingface_sagemaker_tensorflow_distributed_b200.ops import gemm as G, _lib
lib = _lib.load(); lib.hstd_gemm_set_trace.argtypes = [ctypes.c_void_p]; lib.hstd_gemm_set_trace.restype = None
def run(M, N, K, bn, cm, a_t=False, b_t=False, nkb=24):
    A = torch.randn((K, M) if a_t else (M, K), device="cuda").bfloat16(); B = torch.randn((K, N) if b_t else (N, K), device="cuda").bfloat16()
    C = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    for _ in range(3): G.gemm(A, B, a_t=a_t, b_t=b_t, out=C, block_n=bn, cluster_m=cm)
    tr = torch.zeros(512, dtype=torch.int64, device="cuda"); lib.hstd_gemm_set_trace(tr.data_ptr())
    G.gemm(A, B, a_t=a_t, b_t=b_t, out=C, block_n=bn, cluster_m=cm); torch.cuda.synchronize(); lib.hstd_gemm_set_trace(None)
    t = tr.cpu().tolist(); t0 = t[0]
    print(f"--- M={M} N={N} K={K} bn={bn} cluster_m={cm} a_t={a_t} b_t={b_t}: (cycles since CTA start) kb: prod_slot_free | mma_wait_start -> data_landed (+wait) -> issued")
    prev = None
    for kb in range(min(nkb, K // 64)):
        a, b, c, d = (t[2 + 4 * kb + i] - t0 for i in range(4))
        print(f"kb {kb:2d}: {a:6d} | {b:6d} -> {c:6d} (+{c-b:5d}) -> {d:6d}" + (f"   period {c-prev}" if prev is not None else ""))
        prev = c
run(768, 768, 4096, 128, 1)
run(768, 768, 4096, 256, 1)
