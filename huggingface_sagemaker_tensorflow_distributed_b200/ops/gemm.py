"""Python front-end of the tcgen05 GEMM (csrc/gemm_sm100.cu).

``gemm(A, B, ...)`` computes ``C[m,n] = epi(alpha * sum_k A(m,k) B(n,k) + bias[n])`` on bf16
CUDA tensors.  ``a_t`` / ``b_t`` mean the operand is stored transposed (``[K, M]`` /
``[K, N]`` row-major), which the kernel reads MN-major without an explicit transpose:

    forward   Y  = X · Wᵀ          gemm(X, W)
    dgrad     dX = dY · W          gemm(dY, W, b_t=True)
    wgrad     dW = dYᵀ · X         gemm(dY, X, a_t=True, b_t=True)

Replaces the cuBLAS matmuls implied by the reference's ``model.fit`` (scripts/train.py:145).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib

EPI_NONE, EPI_GELU, EPI_ADD, EPI_GELU_BWD = 0, 1, 2, 3

_TUNING = None


def tuned_config(M: int, N: int, K: int, a_t: bool, b_t: bool):
    """(block_n, cluster_m, splits) measured best on B200 by tools/gemm_tune.py for this shape,
    or None (the kernel's own heuristics apply)."""
    global _TUNING
    if _TUNING is None:
        import json
        import os
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "gemm_tuning.json")
        try:
            with open(path) as f:
                _TUNING = json.load(f)
        except (OSError, ValueError):
            _TUNING = {}
    e = _TUNING.get(f"{M},{N},{K},{int(a_t)},{int(b_t)}")
    if e is None:
        return None
    return int(e["block_n"]), int(e["cluster_m"]), int(e["splits"])


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def gemm_raw(args: _lib.GemmArgs, stream=None, splits: int = 1, workspace: Optional[torch.Tensor] = None) -> None:
    lib = _lib.load()
    if splits > 1:
        assert workspace is not None
        rc = lib.hstd_gemm_bf16_splitk(C.byref(args), splits, workspace.data_ptr(),
                                       workspace.numel() * workspace.element_size(), _lib.stream_ptr(stream))
    else:
        rc = lib.hstd_gemm_bf16(C.byref(args), _lib.stream_ptr(stream))
    _lib.check(rc, "hstd_gemm_bf16")


def gemm(A: torch.Tensor, B: torch.Tensor, *, a_t: bool = False, b_t: bool = False,
         bias: Optional[torch.Tensor] = None, epilogue: int = EPI_NONE, aux: Optional[torch.Tensor] = None,
         alpha: float = 1.0, out: Optional[torch.Tensor] = None, block_n: int = 0, splits: int = 1,
         workspace: Optional[torch.Tensor] = None, max_ctas: int = 0, cluster_m: int = 0,
         stream=None) -> torch.Tensor:
    """2-D (or 3-D batched, leading batch dim) bf16 GEMM.  See module docstring."""
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16 and A.is_cuda and B.is_cuda
    batched = A.dim() == 3
    if batched:
        assert B.dim() == 3 and A.shape[0] == B.shape[0]
    assert A.stride(-1) == 1 and B.stride(-1) == 1
    nb = A.shape[0] if batched else 1
    if a_t:
        K, M = A.shape[-2], A.shape[-1]
    else:
        M, K = A.shape[-2], A.shape[-1]
    if b_t:
        Kb, N = B.shape[-2], B.shape[-1]
    else:
        N, Kb = B.shape[-2], B.shape[-1]
    assert K == Kb, (A.shape, B.shape, a_t, b_t)
    if out is None:
        out = torch.empty((nb, M, N) if batched else (M, N), dtype=torch.bfloat16, device=A.device)
    assert out.stride(-1) == 1
    if splits > 1 and workspace is None:
        nbytes = _lib.load().hstd_gemm_splitk_workspace_bytes(M, N, splits)
        workspace = torch.empty(nbytes // 4, dtype=torch.float32, device=A.device)
    g = _lib.GemmArgs()
    g.A, g.B, g.C, g.bias, g.aux = _ptr(A), _ptr(B), _ptr(out), _ptr(bias), _ptr(aux)
    g.lda, g.ldb, g.ldc = A.stride(-2), B.stride(-2), out.stride(-2)
    g.a_bs0 = A.stride(0) if batched else 0
    g.b_bs0 = B.stride(0) if batched else 0
    g.c_bs0 = out.stride(0) if batched else 0
    g.a_bs1 = g.b_bs1 = g.c_bs1 = 0
    g.ldaux = aux.stride(-2) if aux is not None else 0
    g.M, g.N, g.K, g.nb0, g.nb1 = M, N, K, nb, 1
    g.a_mn_major, g.b_mn_major = int(a_t), int(b_t)
    g.epilogue, g.alpha, g.block_n, g.max_ctas, g.cluster_m = epilogue, alpha, block_n, max_ctas, cluster_m
    gemm_raw(g, stream, splits, workspace)
    return out
