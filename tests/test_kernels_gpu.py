"""Row kernels, head/loss and optimizer vs PyTorch fp32 statements of the same math."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def K():
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import kernels
    return kernels


def bf(*shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).bfloat16()


def close(got, want, atol=2e-2, rtol=2e-2):
    got, want = got.float(), want.float()
    err = (got - want).abs()
    ok = err <= atol + rtol * want.abs()
    assert ok.all(), f"max err {err.max().item():.4g} at {err.argmax().item()} (want max {want.abs().max().item():.3g})"


@pytest.mark.parametrize("d", [256, 768, 1024])
def test_residual_layernorm_fwd_bwd(d):
    k = K()
    torch.manual_seed(0)
    M = 520
    a, x, gamma, beta = bf(M, d), bf(M, d), (1 + 0.1 * torch.randn(d, device="cuda")).bfloat16(), bf(d, scale=0.1)
    pre, y = torch.empty_like(a), torch.empty_like(a)
    mean, rstd = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
    k.ln_fwd(a=a, resid=x, gamma=gamma, beta=beta, eps=1e-12, pre=pre, y=y, mean=mean, rstd=rstd)
    pre_ref = (a.float() + x.float()).bfloat16().float().requires_grad_(True)
    y_ref = F.layer_norm(pre_ref, (d,), gamma.float(), beta.float(), 1e-12)
    close(pre, pre_ref)
    close(y, y_ref)
    close(mean, pre_ref.mean(-1), 1e-3, 1e-3)
    dy = bf(M, d)
    gw, gb = gamma.float().requires_grad_(True), beta.float().requires_grad_(True)
    F.layer_norm(pre_ref, (d,), gw, gb, 1e-12).backward(dy.float())
    dpre = torch.empty_like(a)
    dgamma, dbeta, dbias = (torch.empty(d, dtype=torch.bfloat16, device="cuda") for _ in range(3))
    partials = torch.zeros(k.ln_bwd_partials_floats(d), device="cuda")  # accumulator: zero once, left zeroed
    k.ln_bwd(dy=dy, pre=pre, mean=mean, rstd=rstd, gamma=gamma, dpre=dpre, dgamma=dgamma, dbeta=dbeta, dbias=dbias,
             partials=partials)
    close(dpre, pre_ref.grad, 3e-2, 3e-2)
    close(dgamma, gw.grad, 0.3, 3e-2)
    close(dbeta, gb.grad, 0.3, 3e-2)
    close(dbias, dpre.float().sum(0), 0.3, 3e-2)


def test_layernorm_dropout_is_replayed_in_backward():
    k = K()
    torch.manual_seed(0)
    M, d, p = 512, 768, 0.1
    st = k.new_step_state("cuda", seed=1234)
    k.step_prologue(st, 1e-3, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.0)
    a, x = bf(M, d).abs() + 1.0, torch.zeros(M, d, dtype=torch.bfloat16, device="cuda")
    gamma, beta = torch.ones(d, dtype=torch.bfloat16, device="cuda"), torch.zeros(d, dtype=torch.bfloat16, device="cuda")
    pre, y = torch.empty_like(a), torch.empty_like(a)
    mean, rstd = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
    k.ln_fwd(a=a, resid=x, gamma=gamma, beta=beta, eps=1e-12, pre=pre, y=y, mean=mean, rstd=rstd, state=st, p_in=p,
             site_in=7)
    keep = pre != 0
    assert abs(keep.float().mean().item() - (1 - p)) < 0.01
    close(pre[keep], (a.float() / (1 - p))[keep], 5e-2, 2e-2)
    dy = bf(M, d)
    dpre, da = torch.empty_like(a), torch.empty_like(a)
    dg, db, dbias = (torch.empty(d, dtype=torch.bfloat16, device="cuda") for _ in range(3))
    partials = torch.zeros(k.ln_bwd_partials_floats(d), device="cuda")  # accumulator: zero once, left zeroed
    k.ln_bwd(dy=dy, pre=pre, mean=mean, rstd=rstd, gamma=gamma, dpre=dpre, da=da, dgamma=dg, dbeta=db, dbias=dbias,
             partials=partials, state=st, p_in=p, site_in=7)
    assert torch.equal(da != 0, keep & (dpre != 0))  # same mask as forward
    close(da[keep], (dpre.float() / (1 - p))[keep], 1e-2, 2e-2)
    # a new step re-keys the mask
    k.step_prologue(st, 1e-3, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.0)
    pre2 = torch.empty_like(a)
    k.ln_fwd(a=a, resid=x, gamma=gamma, beta=beta, eps=1e-12, pre=pre2, y=y, mean=mean, rstd=rstd, state=st, p_in=p,
             site_in=7)
    assert ((pre2 != 0) != keep).float().mean().item() > 0.05


def test_embedding_layernorm_fwd_and_scatter_bwd():
    k = K()
    torch.manual_seed(0)
    B, S, d, V = 4, 128, 768, 1000
    ids = torch.randint(0, V, (B, S), device="cuda", dtype=torch.int32)
    ids[:, 100:] = 1  # padding: many duplicates
    pos = torch.empty_like(ids)
    k.position_ids(ids, pos, roberta=True, pad_id=1)
    nonpad = (ids != 1).long()
    assert torch.equal(pos.long(), torch.cumsum(nonpad, 1) * nonpad + 1)
    k.position_ids(ids, pos, roberta=False, pad_id=0)
    assert torch.equal(pos[0].long(), torch.arange(S, device="cuda"))
    word, post, type0 = bf(V, d, scale=0.05), bf(512, d, scale=0.05), bf(d, scale=0.05)
    gamma, beta = (1 + 0.1 * torch.randn(d, device="cuda")).bfloat16(), bf(d, scale=0.1)
    y, pre = torch.empty(B * S, d, dtype=torch.bfloat16, device="cuda"), torch.empty(B * S, d, dtype=torch.bfloat16, device="cuda")
    mean, rstd = torch.empty(B * S, device="cuda"), torch.empty(B * S, device="cuda")
    k.ln_fwd(y=y, pre=pre, mean=mean, rstd=rstd, gamma=gamma, beta=beta, eps=1e-12, ids=ids.view(-1),
             pos_ids=pos.view(-1), word=word, pos=post, type0=type0)
    e = word.float()[ids.view(-1).long()] + post.float()[pos.view(-1).long()] + type0.float()
    close(y, F.layer_norm(e.bfloat16().float(), (d,), gamma.float(), beta.float(), 1e-12))
    de = bf(B * S, d, scale=0.01)
    wg = torch.zeros(V, d, dtype=torch.bfloat16, device="cuda")
    pg = torch.zeros(512, d, dtype=torch.bfloat16, device="cuda")
    k.embed_bwd(de, ids.view(-1), pos.view(-1), wg, pg)
    wref = torch.zeros(V, d, device="cuda").index_add_(0, ids.view(-1).long(), de.float())
    pref = torch.zeros(512, d, device="cuda").index_add_(0, pos.view(-1).long(), de.float())
    close(wg, wref, 2e-2, 5e-2)  # bf16 atomics
    close(pg, pref, 2e-3, 3e-2)


@pytest.mark.parametrize("N", [768, 2304])
def test_colsum(N):
    k = K()
    x = bf(4096, N)
    out = torch.empty(N, dtype=torch.bfloat16, device="cuda")
    k.colsum(x, out, torch.empty(k.colsum_partials_floats(N), device="cuda"))
    close(out, x.float().sum(0), 0.5, 1e-2)


@pytest.mark.parametrize("Sk", [128, 512])
def test_softmax_fwd_bwd_with_padding_mask(Sk):
    k = K()
    torch.manual_seed(0)
    B, H, Sq = 2, 3, 128
    s = bf(B, H, Sq, Sk, scale=2.0)
    mask = torch.ones(B, Sk, dtype=torch.int32, device="cuda")
    mask[1, Sk // 2:] = 0
    sf = s.float().requires_grad_(True)
    add = torch.zeros(B, 1, 1, Sk, device="cuda").masked_fill(mask[:, None, None, :] == 0, float("-inf"))
    pref = torch.softmax(sf + add, -1)
    P = s.clone()
    k.softmax_fwd(P, mask)
    close(P, pref, 2e-3, 2e-2)
    assert (P[1, :, :, Sk // 2:] == 0).all()
    dP = bf(B, H, Sq, Sk)
    pref.backward(dP.float())
    dS = dP.clone()
    k.softmax_bwd(P, dS)
    close(dS, sf.grad, 5e-3, 3e-2)


def test_softmax_dropout_consistency():
    k = K()
    st = k.new_step_state("cuda", seed=99)
    k.step_prologue(st, 1e-3, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.0)
    B, H, Sq, Sk, p = 2, 2, 128, 512, 0.1
    s = bf(B, H, Sq, Sk)
    mask = torch.ones(B, Sk, dtype=torch.int32, device="cuda")
    P, Pd = s.clone(), torch.empty_like(s)
    k.softmax_fwd(P, mask, pd=Pd, p_drop=p, site=3, state=st)
    keep = Pd != 0
    assert abs(keep.float().mean().item() - (1 - p)) < 0.01
    close(Pd[keep], (P.float() / (1 - p))[keep], 1e-3, 2e-2)
    dPd = bf(B, H, Sq, Sk)
    dS = dPd.clone()
    k.softmax_bwd(P, dS, p_drop=p, site=3, state=st)
    dP = dPd.float() * keep.float() / (1 - p)
    Pf = P.float()
    close(dS, Pf * (dP - (dP * Pf).sum(-1, keepdim=True)), 5e-3, 3e-2)


@pytest.mark.parametrize("kind", ["distilbert", "bert", "roberta"])
def test_head_loss_fwd_bwd(kind):
    k = K()
    torch.manual_seed(0)
    B, S, d, C = 8, 16, 768, 2
    x = bf(B * S, d)
    w1, b1, w2, b2 = bf(d, d, scale=0.05), bf(d, scale=0.1), bf(C, d, scale=0.2), bf(C, scale=0.1)
    labels = torch.tensor([0, 1, 1, 0, -1, 1, 0, 1], dtype=torch.int32, device="cuda")
    st = k.new_step_state("cuda", seed=5)
    logits, dlogits = torch.empty(B, C, device="cuda"), torch.empty(B, C, device="cuda")
    h, dz = torch.empty(B, d, device="cuda"), torch.empty(B, d, device="cuda")
    dw1, db1, dw2, db2 = torch.empty_like(w1), torch.empty_like(b1), torch.empty_like(w2), torch.empty_like(b2)
    dx = torch.full((B * S, d), 7.0, dtype=torch.bfloat16, device="cuda")
    args = k.make_head_args(x=x, w1=w1, b1=b1, w2=w2, b2=b2, labels=labels, logits=logits, dlogits=dlogits, h=h, dz=dz,
                            state=st, B=B, S=S, d=d, num_labels=C, kind=kind, p_drop=0.0, site0=1, site1=2, dw1=dw1,
                            db1=db1, dw2=dw2, db2=db2, dx=dx)
    k.head_fwd(args, train=True)
    k.head_bwd(args)
    xc = x.view(B, S, d)[:, 0].float().requires_grad_(True)
    W1, B1, W2, B2 = (t.float().requires_grad_(True) for t in (w1, b1, w2, b2))
    z = xc @ W1.t() + B1
    hh = torch.relu(z) if kind == "distilbert" else torch.tanh(z)
    lg = hh @ W2.t() + B2
    valid = labels >= 0
    loss_sum = F.cross_entropy(lg[valid], labels[valid].long(), reduction="sum")
    (loss_sum / valid.sum()).backward()
    close(logits, lg, 2e-3, 2e-3)
    s = k.read_step_state(st)
    assert s.metrics[0] == pytest.approx(loss_sum.item(), rel=2e-3)
    assert s.metrics[1] == (lg.argmax(-1)[valid] == labels[valid]).sum().item() and s.metrics[2] == 7
    close(dw2, W2.grad, 2e-3, 2e-2)
    close(db2, B2.grad, 2e-3, 2e-2)
    close(dw1, W1.grad, 2e-3, 3e-2)
    close(db1, B1.grad, 2e-3, 2e-2)
    close(dx.view(B, S, d)[:, 0], xc.grad, 2e-3, 2e-2)
    assert (dx.view(B, S, d)[:, 1:] == 0).all()


@pytest.mark.parametrize("kind", ["distilbert", "bert"])
def test_head_dropout_mask_is_shared_between_forward_and_backward(kind):
    """Dropout on the hidden activation: the forward leaves h*m in h, so the mask can be read
    back (m = 0 or 1/keep) and every output checked against autograd with that same mask."""
    k = K()
    torch.manual_seed(1)
    B, S, d, C, p = 8, 4, 768, 3, 0.2
    x = bf(B * S, d)
    w1, b1, w2, b2 = bf(d, d, scale=0.05), bf(d, scale=0.1), bf(C, d, scale=0.2), bf(C, scale=0.1)
    labels = torch.tensor([0, 1, 2, 0, 2, 1, -1, 1], dtype=torch.int32, device="cuda")
    st = k.new_step_state("cuda", seed=9)
    logits, dlogits = torch.empty(B, C, device="cuda"), torch.empty(B, C, device="cuda")
    h, dz = torch.empty(B, d, device="cuda"), torch.empty(B, d, device="cuda")
    dw1, db1, dw2, db2 = torch.empty_like(w1), torch.empty_like(b1), torch.empty_like(w2), torch.empty_like(b2)
    dx = torch.empty((B * S, d), dtype=torch.bfloat16, device="cuda")
    args = k.make_head_args(x=x, w1=w1, b1=b1, w2=w2, b2=b2, labels=labels, logits=logits, dlogits=dlogits, h=h, dz=dz,
                            state=st, B=B, S=S, d=d, num_labels=C, kind=kind, p_drop=p, site0=1, site1=2, dw1=dw1,
                            db1=db1, dw2=dw2, db2=db2, dx=dx)
    k.head_fwd(args, train=True)
    hd = h.clone()
    k.head_bwd(args)
    xc = x.view(B, S, d)[:, 0].float().requires_grad_(True)
    W1, B1, W2, B2 = (t.float().requires_grad_(True) for t in (w1, b1, w2, b2))
    z = xc @ W1.t() + B1
    hh = torch.relu(z) if kind == "distilbert" else torch.tanh(z)
    keep_q = 1.0 - round(p * 256) / 256.0
    big = hh.detach().abs() > 1e-3
    m = torch.where(hd.abs() > 0, torch.full_like(hd, 1.0 / keep_q), torch.zeros_like(hd))
    # where the activation is clearly non-zero the stored value is exactly h/keep or 0
    ratio = (hd / hh.detach())[big]
    assert ((ratio.abs() < 1e-6) | ((ratio - 1.0 / keep_q).abs() < 2e-2)).all()
    drop_rate = 1.0 - (ratio.abs() > 1e-6).float().mean().item()
    assert abs(drop_rate - (1.0 - keep_q)) < 0.03
    m = torch.where(big, m, torch.full_like(m, 1.0 / keep_q))  # relu zeros: the mask is unobservable and irrelevant
    lg = (hh * m) @ W2.t() + B2
    valid = labels >= 0
    (F.cross_entropy(lg[valid], labels[valid].long(), reduction="sum") / valid.sum()).backward()
    if kind == "distilbert":
        close(logits, lg, 2e-3, 2e-3)
        close(dw2, W2.grad, 2e-3, 2e-2)
        close(dw1, W1.grad, 2e-3, 3e-2)
        close(db1, B1.grad, 2e-3, 2e-2)
        close(dx.view(B, S, d)[:, 0], xc.grad, 2e-3, 2e-2)
    else:  # tanh: tiny activations were given mask 1/keep above, so allow for the few that were dropped
        close(logits, lg, 5e-3, 5e-3)
        close(dw2, W2.grad, 5e-3, 5e-2)
    # a second forward + backward of the same step regenerates the same mask
    k.head_fwd(args, train=True)
    assert torch.equal(h, hd)


def test_adamw_flat_matches_keras_formula_and_prologue_schedule():
    k = K()
    from huggingface_sagemaker_tensorflow_distributed_b200.optim import KerasAdamW, LRSchedule
    torch.manual_seed(0)
    n = 4096 + 128
    w0 = torch.randn(n, device="cuda")
    master, m, v = w0.clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    wb = torch.empty(n, dtype=torch.bfloat16, device="cuda")
    st = k.new_step_state("cuda", seed=1)
    p = torch.nn.Parameter(w0.clone())
    sched = LRSchedule(1e-2, "linear", total_steps=10, warmup_steps=2)
    opt = KerasAdamW([p], learning_rate=1e-2, weight_decay=0.01, schedule=sched)
    for step in range(5):
        g = (torch.randn(n, device="cuda") * (1e-6 if step == 2 else 0.1)).bfloat16()
        k.step_prologue(st, 1e-2, "linear", 10, 2, 0.9, 0.999, 1e-7, 0.01)
        k.adamw_flat(st, g, master, m, v, wb)
        p.grad = g.float()
        opt.step()
        s = k.read_step_state(st)
        assert s.step == step + 1 and s.lr == pytest.approx(sched(step), rel=1e-6)
    assert torch.allclose(master, p.data, atol=1e-5, rtol=1e-5)
    assert torch.equal(wb, master.bfloat16())
