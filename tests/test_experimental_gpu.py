"""Kernels under csrc/experimental/ — only when the library was built with ``--experimental``
(``python -m huggingface_sagemaker_tensorflow_distributed_b200.build --experimental --watchdog``);
skipped otherwise, so the default test run never touches unvalidated code."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _pp():
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import _lib

    lib = _lib.load()
    if not hasattr(lib, "hstd_attn_fwd_pp"):
        pytest.skip("library built without --experimental")
    fn = lib.hstd_attn_fwd_pp
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float,
                   C.c_uint32, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    return fn, _lib


@pytest.mark.parametrize("S", [256, 512])
@pytest.mark.parametrize("p_drop", [0.0, 0.1])
def test_pingpong_forward_matches_the_kernel_in_the_tree(S, p_drop):
    """Same inputs, same dropout stream: context and log-sum-exp must agree with hstd_attn_fwd to bf16
    rounding (the two kernels differ in summation order and in the reference maximum only)."""
    fn, _lib = _pp()
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import kernels as K

    torch.manual_seed(0)
    B, H = 2, 3
    d = H * 64
    qkv = torch.randn(B * S, 3 * d, device="cuda").bfloat16()
    qkv[:, :d] *= 3.0  # large scores in places: exercises the lazy rescale of the reference maximum
    mask = torch.ones(B, S, dtype=torch.int32, device="cuda")
    mask[1, S - 37:] = 0
    st = K.new_step_state("cuda", seed=3)
    K.step_prologue(st, 1e-3, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.0)
    ctx0, ctx1 = (torch.full((B * S, d), float("nan"), dtype=torch.bfloat16, device="cuda") for _ in range(2))
    lse0, lse1 = (torch.full((B, H, S), float("nan"), device="cuda") for _ in range(2))
    K.attn_fwd(qkv, ctx0, lse0, mask, B, H, S, 0.125, p_drop=p_drop, site=5, state=st)
    rc = fn(qkv.data_ptr(), ctx1.data_ptr(), lse1.data_ptr(), mask.data_ptr(), B, H, S, 0.125, p_drop, 5, st.data_ptr(),
            torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "hstd_attn_fwd_pp")
    torch.cuda.synchronize()
    assert torch.isfinite(ctx1.float()).all() and torch.isfinite(lse1).all()
    assert torch.allclose(lse1, lse0, atol=2e-3, rtol=1e-4)
    err = (ctx1.float() - ctx0.float()).abs().max().item()
    assert err <= 0.02 * ctx0.float().abs().max().item() + 2e-3, err


@pytest.mark.parametrize("S", [128, 256, 512])
@pytest.mark.parametrize("p_drop", [0.0, 0.1])
def test_sixteen_warp_backward_matches_the_kernel_in_the_tree(S, p_drop):
    """Same schedule and barriers, softmax side spread over 16 warps: dK, dV and the summed dQ partials must
    agree with hstd_attn_bwd (identical dropout stream; only the staging granularity differs)."""
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import _lib
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import kernels as K

    lib = _lib.load()
    if not hasattr(lib, "hstd_attn_bwd_w16_core"):
        pytest.skip("library built without --experimental")
    fn = lib.hstd_attn_bwd_w16_core
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                   C.c_float, C.c_float, C.c_uint32, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    torch.manual_seed(0)
    B, H = 2, 3
    d = H * 64
    qkv = torch.randn(B * S, 3 * d, device="cuda").bfloat16()
    mask = torch.ones(B, S, dtype=torch.int32, device="cuda")
    mask[1, S - 37:] = 0
    dO = (torch.randn(B * S, d, device="cuda") * 0.5).bfloat16()
    st = K.new_step_state("cuda", seed=3)
    K.step_prologue(st, 1e-3, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.0)
    ctx = torch.empty(B * S, d, dtype=torch.bfloat16, device="cuda")
    lse = torch.empty(B, H, S, device="cuda")
    K.attn_fwd(qkv, ctx, lse, mask, B, H, S, 0.125, p_drop=p_drop, site=5, state=st)
    ws = torch.empty(K.attn_bwd_workspace_bytes(B, H, S), dtype=torch.uint8, device="cuda")
    ref = torch.full((B * S, 3 * d), float("nan"), dtype=torch.bfloat16, device="cuda")
    K.attn_bwd(qkv, ctx, dO, lse, mask, ref, ws, B, H, S, 0.125, p_drop=p_drop, site=5, state=st)  # also leaves delta in ws
    got = torch.full((B * S, 3 * d), float("nan"), dtype=torch.bfloat16, device="cuda")
    rc = fn(qkv.data_ptr(), dO.data_ptr(), lse.data_ptr(), mask.data_ptr(), got.data_ptr(), ws.data_ptr(), B, H, S, 0.125,
            p_drop, 5, st.data_ptr(), torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "hstd_attn_bwd_w16_core")
    torch.cuda.synchronize()
    parts = ws[B * H * S * 4:].view(torch.bfloat16).view(S // 128, B * S, d).float().sum(0)
    for name, a, b in (("dq", parts, ref[:, :d].float()), ("dk", got[:, d:2 * d].float(), ref[:, d:2 * d].float()),
                       ("dv", got[:, 2 * d:].float(), ref[:, 2 * d:].float())):
        assert torch.isfinite(a).all(), name
        assert (a - b).norm() <= 0.01 * b.norm() + 1e-3, (name, ((a - b).norm() / b.norm()).item())
