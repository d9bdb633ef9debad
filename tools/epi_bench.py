import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from huggingface_sagemaker_tensorflow_distributed_b200.ops import gemm as G
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gemm_tune import graph_time
M, N, K = 4096, 3072, 768
X = torch.randn(M, K, device="cuda").bfloat16(); W = torch.randn(N, K, device="cuda").bfloat16() * 0.05
b = torch.randn(N, device="cuda").bfloat16(); C = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
aux = torch.empty_like(C); R = torch.randn(M, N, device="cuda").bfloat16()
for name, kw in [("plain", {}), ("bias", dict(bias=b)), ("gelu (1 store)", dict(bias=b, epilogue=G.EPI_GELU)),
                 ("gelu + pre-act store", dict(bias=b, epilogue=G.EPI_GELU, aux=aux)),
                 ("add residual (1 load)", dict(epilogue=G.EPI_ADD, aux=R)),
                 ("gelu_bwd (1 load)", dict(epilogue=G.EPI_GELU_BWD, aux=R))]:
    us = graph_time(lambda: G.gemm(X, W, out=C, **kw))
    print(f"{name:28s} {us:6.1f} us  {2*M*N*K/us/1e6:6.0f} TFLOP/s")
