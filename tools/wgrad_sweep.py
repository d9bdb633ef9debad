"""All (block_n, 1-SM/2-SM, split-K) timings for the DistilBERT/BERT-base weight-gradient shapes."""
import itertools, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from huggingface_sagemaker_tensorflow_distributed_b200.ops import gemm as G
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gemm_tune import graph_time
for (M, N, K) in [(3072, 768, 4096), (768, 3072, 4096), (2304, 768, 4096), (768, 768, 4096)]:
    A = torch.randn(K, M, device="cuda").bfloat16(); B = torch.randn(K, N, device="cuda").bfloat16()
    C = torch.empty(M, N, dtype=torch.bfloat16, device="cuda"); ws = torch.empty(M * N * 8, device="cuda")
    row = []
    for bn, mc, sp in itertools.product((128, 256), (1, 2), (1, 2, 3, 4, 8)):
        us = graph_time(lambda: G.gemm(A, B, a_t=True, b_t=True, out=C, block_n=bn, cluster_m=(1 if mc == 1 else 0), splits=sp, workspace=ws))
        row.append(f"bn{bn}/{mc}sm/s{sp}={us:.1f}")
    print((M, N, K), " ".join(row))
