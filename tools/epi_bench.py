import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from huggingface_sagemaker_tensorflow_distributed_b200.ops import gemm as G
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gemm_tune import graph_time
def run(M, N, K, **kw):
    X = torch.randn(M, K, device="cuda").bfloat16(); W = torch.randn(N, K, device="cuda").bfloat16() * 0.05
    C = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    return graph_time(lambda: G.gemm(X, W, out=C, **kw))
for (M, N) in [(4096, 2304), (4096, 768)]:
    for K in (64, 768, 1536, 3072, 6144):
        a = run(M, N, K, block_n=256, cluster_m=1)
        b = run(M, N, K, block_n=256, cluster_m=1, epilogue=100)
        c = run(M, N, K, block_n=256, cluster_m=0)
        d = run(M, N, K, block_n=256, cluster_m=0, epilogue=100)
        Cr = torch.empty(M, N, dtype=torch.bfloat16, device="cuda"); X = torch.randn(M, K, device="cuda").bfloat16(); W = torch.randn(N, K, device="cuda").bfloat16()
        r = graph_time(lambda: torch.matmul(X, W.t(), out=Cr))
        print(f"M={M} N={N} K={K:5d}: 1SM {a:6.1f} (no-store {b:6.1f})  2SM {c:6.1f} (no-store {d:6.1f})  cuBLAS {r:6.1f} us")
