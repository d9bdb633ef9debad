"""Fused tcgen05 attention vs a PyTorch fp32 statement of the same op (and vs the unfused
softmax kernel for the dropout mask)."""

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(qkv, mask, B, S, H, dO=None, drop_keep=None, p=0.0):
    d = H * 64
    x = qkv.float().detach().requires_grad_(True)
    q, k, v = (x.view(B, S, 3, H, 64)[:, :, i].transpose(1, 2) for i in range(3))
    s = q @ k.transpose(-1, -2) / 8.0
    s = s + torch.zeros(B, 1, 1, S, device=x.device).masked_fill(mask[:, None, None, :] == 0, float("-inf"))
    P = torch.softmax(s, -1)
    Pd = P if drop_keep is None else P * drop_keep / (1 - p)
    ctx = (Pd @ v).transpose(1, 2).reshape(B * S, d)
    lse = torch.logsumexp(s, -1)
    if dO is not None:
        ctx.backward(dO.float())
        return ctx.detach(), lse.detach(), x.grad
    return ctx.detach(), lse.detach(), None


def _close(got, want, tol=0.03):
    err = (got.float() - want).abs().max().item()
    assert err <= tol * want.abs().max().item() + 2e-3, (err, want.abs().max().item())


@pytest.mark.parametrize("S", [128, 256, 512])
def test_fused_attention_fwd_bwd_matches_reference(S):
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import kernels as K
    torch.manual_seed(0)
    B, H = 2, 3
    d = H * 64
    qkv = (torch.randn(B * S, 3 * d, device="cuda")).bfloat16()
    mask = torch.ones(B, S, dtype=torch.int32, device="cuda")
    mask[1, S - 37:] = 0
    dO = (torch.randn(B * S, d, device="cuda") * 0.5).bfloat16()
    ctx = torch.empty(B * S, d, dtype=torch.bfloat16, device="cuda")
    lse = torch.empty(B, H, S, device="cuda")
    K.attn_fwd(qkv, ctx, lse, mask, B, H, S, 0.125)
    rctx, rlse, rgrad = _ref(qkv, mask, B, S, H, dO)
    _close(ctx, rctx)
    assert torch.allclose(lse, rlse, atol=2e-2, rtol=1e-3)
    dqkv = torch.full((B * S, 3 * d), float("nan"), dtype=torch.bfloat16, device="cuda")
    ws = torch.empty(K.attn_bwd_workspace_bytes(B, H, S), dtype=torch.uint8, device="cuda")
    K.attn_bwd(qkv, ctx, dO, lse, mask, dqkv, ws, B, H, S, 0.125)
    for name, sl in (("dq", slice(0, d)), ("dk", slice(d, 2 * d)), ("dv", slice(2 * d, 3 * d))):
        got, want = dqkv[:, sl].float(), rgrad[:, sl]
        assert torch.isfinite(got).all(), name
        rel = (got - want).norm() / want.norm()
        assert rel < 0.03, (name, rel.item())


def test_fused_attention_dropout_uses_the_same_mask_as_the_softmax_kernel():
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import kernels as K
    torch.manual_seed(1)
    B, H, S, p = 2, 2, 256, 0.1
    d = H * 64
    st = K.new_step_state("cuda", seed=77)
    K.step_prologue(st, 1e-3, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.0)
    qkv = torch.randn(B * S, 3 * d, device="cuda").bfloat16()
    mask = torch.ones(B, S, dtype=torch.int32, device="cuda")
    # recover the keep mask from the unfused softmax kernel (same site / seed / step)
    scores = torch.zeros(B, H, S, S, dtype=torch.bfloat16, device="cuda")
    Pd = torch.empty_like(scores)
    K.softmax_fwd(scores, mask, pd=Pd, p_drop=p, site=5, state=st)
    keep = (Pd != 0).float()
    assert abs(keep.mean().item() - (1 - p)) < 0.01
    dO = (torch.randn(B * S, d, device="cuda") * 0.5).bfloat16()
    ctx = torch.empty(B * S, d, dtype=torch.bfloat16, device="cuda")
    lse = torch.empty(B, H, S, device="cuda")
    K.attn_fwd(qkv, ctx, lse, mask, B, H, S, 0.125, p_drop=p, site=5, state=st)
    p_eff = round(p * 256) / 256.0
    rctx, _, rgrad = _ref(qkv, mask, B, S, H, dO, drop_keep=keep, p=p_eff)
    _close(ctx, rctx)
    dqkv = torch.empty(B * S, 3 * d, dtype=torch.bfloat16, device="cuda")
    ws = torch.empty(K.attn_bwd_workspace_bytes(B, H, S), dtype=torch.uint8, device="cuda")
    K.attn_bwd(qkv, ctx, dO, lse, mask, dqkv, ws, B, H, S, 0.125, p_drop=p, site=5, state=st)
    rel = (dqkv.float() - rgrad).norm() / rgrad.norm()
    assert rel < 0.03, rel.item()
