"""tcgen05 GEMM vs a PyTorch fp32 reference of the same product (bf16 inputs, fp32 math)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(A, B, a_t, b_t):
    Af = A.float().transpose(-1, -2) if a_t else A.float()
    Bf = B.float() if b_t else B.float().transpose(-1, -2)
    return Af @ Bf


def _mk(shape, scale=0.5):
    return (torch.randn(*shape, device="cuda") * scale).bfloat16()


def _close(got, want, tol=0.02):
    err = (got.float() - want).abs().max().item()
    assert err <= tol * want.abs().max().item() + 1e-2, err


@pytest.mark.parametrize("a_t,b_t", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("bn", [64, 128, 192, 256])
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (384, 768, 192), (200, 328, 136)])
def test_all_operand_layouts(a_t, b_t, bn, M, N, K):
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import gemm as G
    torch.manual_seed(0)
    A = _mk((K, M) if a_t else (M, K))
    B = _mk((K, N) if b_t else (N, K))
    _close(G.gemm(A, B, a_t=a_t, b_t=b_t, block_n=bn), _ref(A, B, a_t, b_t))


def test_model_shapes_persistent_multi_tile():
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import gemm as G
    torch.manual_seed(1)
    X, W = _mk((4096, 768)), _mk((2304, 768))
    _close(G.gemm(X, W), _ref(X, W, False, False))
    dY, W2 = _mk((4096, 768)), _mk((768, 3072))
    _close(G.gemm(dY, W2, b_t=True), _ref(dY, W2, False, True))  # dgrad
    Xf = _mk((4096, 3072))
    _close(G.gemm(dY, Xf, a_t=True, b_t=True), _ref(dY, Xf, True, True))  # wgrad


def test_batched_attention_products():
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import gemm as G
    torch.manual_seed(2)
    Q, K = _mk((24, 512, 64)), _mk((24, 512, 64))
    _close(G.gemm(Q, K, alpha=0.125), 0.125 * _ref(Q, K, False, False))
    P, V = _mk((24, 512, 512), 0.1), _mk((24, 512, 64))
    _close(G.gemm(P, V, b_t=True), _ref(P, V, False, True))
    _close(G.gemm(P, V, a_t=True, b_t=True), _ref(P, V, True, True))


def test_epilogues_bias_gelu_add_gelubwd():
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import gemm as G
    torch.manual_seed(3)
    X, W, b = _mk((512, 256)), _mk((768, 256), 0.1), _mk((768,))
    z = _ref(X, W, False, False) + b.float()
    _close(G.gemm(X, W, bias=b), z)
    pre = torch.empty(512, 768, dtype=torch.bfloat16, device="cuda")
    out = G.gemm(X, W, bias=b, epilogue=G.EPI_GELU, aux=pre)
    _close(pre, z)
    _close(out, torch.nn.functional.gelu(z))
    R = _mk((512, 768))
    _close(G.gemm(X, W, bias=b, epilogue=G.EPI_ADD, aux=R), z + R.float())
    u = _mk((512, 768), 1.5)
    uf = u.float().requires_grad_(True)
    torch.nn.functional.gelu(uf).sum().backward()
    _close(G.gemm(X, W, epilogue=G.EPI_GELU_BWD, aux=u), _ref(X, W, False, False) * uf.grad)


@pytest.mark.parametrize("cluster_m", [0, 1])
def test_192_wide_tiles_multi_tile_epilogues_and_split(cluster_m):
    """BN = 192: 128 tiles for the N = 768 projections (one per SM instead of 96), accumulators
    at TMEM columns 0 / 256, store tiles split 128 + 64 between the warps of a lane quarter."""
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import gemm as G
    torch.manual_seed(5)
    X, W, b = _mk((4096, 3072), 0.5), _mk((768, 3072), 0.05), _mk((768,))
    z = _ref(X, W, False, False) + b.float()
    _close(G.gemm(X, W, bias=b, block_n=192, cluster_m=cluster_m), z)
    R = _mk((4096, 768))
    _close(G.gemm(X, W, bias=b, epilogue=G.EPI_ADD, aux=R, block_n=192, cluster_m=cluster_m), z + R.float())
    pre = torch.empty(4096, 768, dtype=torch.bfloat16, device="cuda")
    out = G.gemm(X, W, bias=b, epilogue=G.EPI_GELU, aux=pre, block_n=192, cluster_m=cluster_m)
    _close(pre, z)
    _close(out, torch.nn.functional.gelu(z))
    dY, W2 = _mk((4096, 2304), 0.5), _mk((2304, 768), 0.05)
    _close(G.gemm(dY, W2, b_t=True, block_n=192, cluster_m=cluster_m), _ref(dY, W2, False, True))  # dgrad (MN-major B)
    Xs = _mk((4096, 768))
    _close(G.gemm(dY, Xs, a_t=True, b_t=True, block_n=192, cluster_m=cluster_m), _ref(dY, Xs, True, True))  # wgrad
    _close(G.gemm(dY, Xs, a_t=True, b_t=True, block_n=192, splits=4, cluster_m=cluster_m), _ref(dY, Xs, True, True))
    # N not a multiple of 192: the last tile column is clipped
    W3 = _mk((1024, 3072), 0.05)
    _close(G.gemm(X, W3, block_n=192, cluster_m=cluster_m), _ref(X, W3, False, False))


@pytest.mark.parametrize("splits", [2, 4, 7])
def test_split_k_wgrad(splits):
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import gemm as G
    torch.manual_seed(4)
    dY, X = _mk((4096, 768)), _mk((4096, 768))
    _close(G.gemm(dY, X, a_t=True, b_t=True, splits=splits, block_n=128), _ref(dY, X, True, True))
