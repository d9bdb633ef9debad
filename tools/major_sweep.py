"""Same GEMM shape with the four operand-major combinations (is MN-major A or B the slow one?)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from huggingface_sagemaker_tensorflow_distributed_b200.ops import gemm as G
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gemm_tune import graph_time
for (M, N, K) in [(3072, 768, 4096), (768, 768, 4096)]:
    for bn in (128, 256):
        out = []
        for a_t in (False, True):
            for b_t in (False, True):
                A = torch.randn((K, M) if a_t else (M, K), device="cuda").bfloat16()
                B = torch.randn((K, N) if b_t else (N, K), device="cuda").bfloat16()
                C = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
                us = graph_time(lambda: G.gemm(A, B, a_t=a_t, b_t=b_t, out=C, block_n=bn, cluster_m=1))
                out.append(f"a_t={int(a_t)} b_t={int(b_t)}: {us:.1f}")
        print((M, N, K), f"bn={bn}", " | ".join(out))
