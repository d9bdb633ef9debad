"""The fused all-reduce kernels with every rank emulated on ONE GPU (SURVEY.md §4 collective tier, for boxes
with fewer GPUs than ranks): one cooperative grid runs the blocks of all ranks — same per-block code as the
multi-GPU kernels, including the peer-memory flag barriers, the fused Keras-AdamW and the row-sparse
embedding path — against a gathered fp32 reference.  (Separate launches that wait for one another must
not share a GPU, so the emulation is a single kernel.)"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref_adam(s, g, w, m, v):
    m.mul_(s.beta1).add_(g, alpha=1 - s.beta1)
    v.mul_(s.beta2).addcmul_(g, g, value=1 - s.beta2)
    w.mul_(s.decay).addcdiv_(m, v.sqrt() + s.eps, value=-s.lr_t)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("mode_name", ["two_shot", "one_shot"])
def test_plain_allreduce_and_fused_adamw(world, mode_name):
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import kernels as K
    from huggingface_sagemaker_tensorflow_distributed_b200.parallel import p2p

    mode = {"two_shot": p2p.AR_TWO_SHOT, "one_shot": p2p.AR_ONE_SHOT}[mode_name]
    dev = torch.device("cuda")
    sizes = [64 * world, 8192, (1 << 18) + 128 * world]
    starts = [0, sizes[0], sizes[0] + sizes[1]]
    total = sum(sizes)
    emu = p2p.EmulatedP2P(world, total, len(sizes), dev)
    # ---- plain averaging all-reduce, three rounds (exercises the epoch protocol), value + cross-rank bit equality
    for rep in range(3):
        for r in range(world):
            torch.manual_seed(100 * rep + r)
            emu.g16[r].copy_((torch.randn(total, device=dev) * (1 + r)).bfloat16())
            emu.w16[r].zero_()
        want = torch.stack([g.float() for g in emu.g16]).sum(0) / world
        for bi, (st, n) in enumerate(zip(starts, sizes)):
            emu.allreduce([emu.args(r, bi, st, n, mode, False) for r in range(world)])
        torch.cuda.synchronize()
        emu.check_errors()
        got = emu.w16 if mode == p2p.AR_ONE_SHOT else emu.g16
        assert all(torch.equal(got[0], g) for g in got), "ranks disagree"
        err = (got[0].float() - want).abs().max().item()
        assert err <= 2 ** -8 * want.abs().max().item() + 1e-6, err  # one bf16 rounding of the fp32 average
    # ---- fused AdamW: sharded state (two-shot) / replicated state (one-shot) vs gathered fp32 reference
    torch.manual_seed(7)
    w0 = torch.randn(total, device=dev)
    rm, rmm, rv = w0.clone(), torch.zeros(total, device=dev), torch.zeros(total, device=dev)

    def owned(r, st, n):
        return (st, n) if mode == p2p.AR_ONE_SHOT else (st + r * (n // world), n // world)

    state = [[(w0[a:a + k].clone(), torch.zeros(k, device=dev), torch.zeros(k, device=dev))
              for a, k in (owned(r, st, n) for st, n in zip(starts, sizes))] for r in range(world)]
    for r in range(world):
        emu.w16[r].copy_(w0.bfloat16())
    sst = K.new_step_state(dev, seed=1)
    for step in range(3):
        for r in range(world):
            torch.manual_seed(1000 + 10 * step + r)
            emu.g16[r].copy_((torch.randn(total, device=dev) * 0.1).bfloat16())
        avg = torch.stack([g.float() for g in emu.g16]).sum(0) / world
        K.step_prologue(sst, 1e-2, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.01)
        for bi, (st, n) in enumerate(zip(starts, sizes)):
            if mode == p2p.AR_TWO_SHOT and step == 1:  # the split schedule: reduce-only launch, update-only launch
                emu.allreduce([emu.args(r, bi, st, n, mode, True, sst, *state[r][bi], phase=1) for r in range(world)])
                emu.allreduce([emu.args(r, bi, st, n, mode, True, sst, *state[r][bi], phase=2, end_barrier=True)
                               for r in range(world)])
            else:
                emu.allreduce([emu.args(r, bi, st, n, mode, True, sst, *state[r][bi]) for r in range(world)])
        torch.cuda.synchronize()
        # the sharded form hands the averaged gradient from its reduce pass to its optimizer pass in bf16
        _ref_adam(K.read_step_state(sst), avg if mode == p2p.AR_ONE_SHOT else avg.bfloat16().float(), rm, rmm, rv)
    emu.check_errors()
    assert all(torch.equal(emu.w16[0], w) for w in emu.w16), "weights differ across ranks"
    tol = 1e-5 if mode == p2p.AR_ONE_SHOT else 2e-3  # (a different summation order can flip a bf16 rounding)
    for bi, (st, n) in enumerate(zip(starts, sizes)):
        for r in range(world):
            a, k = owned(r, st, n)
            assert torch.allclose(state[r][bi][0], rm[a:a + k], atol=tol, rtol=tol), (bi, r)
    assert (emu.w16[0].float() - rm).abs().max().item() <= 2 ** -8 * rm.abs().max().item() + 2e-3


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("schedule", ["one_launch", "split"])
def test_row_sparse_embedding_update_equals_the_dense_update(world, schedule):
    """one_launch: mark -> idle rows (zero-gradient Adam step, no reduction) -> live rows (reduce + AdamW);
    split: mark -> reduce of the live rows -> one update pass over all rows (live: reduced gradient, idle: zero).
    Either must leave master / m / v / bf16 weights where one dense reduce + AdamW over the whole table leaves
    them, with gradient rows of earlier steps still lying around in the arena (they are never zeroed densely)."""
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import kernels as K
    from huggingface_sagemaker_tensorflow_distributed_b200.parallel import p2p

    dev = torch.device("cuda")
    d, dense_rows, vocab, n_ids = 256, 6, 1000, 96
    rows = dense_rows + vocab
    while (rows * d) % (8 * world):
        rows += 1
    vocab = rows - dense_rows
    total = rows * d
    emu = p2p.EmulatedP2P(world, total, 1, dev, row_elems=d, rows=rows, blocks=6)
    torch.manual_seed(3)
    w0 = torch.randn(total, device=dev)
    n = total // world
    state = [(w0[r * n:(r + 1) * n].clone(), torch.zeros(n, device=dev), torch.zeros(n, device=dev)) for r in range(world)]
    for r in range(world):
        emu.w16[r].copy_(w0.bfloat16())
        emu.g16[r].fill_(float("nan"))  # stale rows must never be read
    rm, rmm, rv = w0.clone(), torch.zeros(total, device=dev), torch.zeros(total, device=dev)
    sst = K.new_step_state(dev, seed=1)
    for step in range(4):
        K.step_prologue(sst, 1e-2, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.0)
        g = torch.Generator(device="cpu").manual_seed(50 + step)
        ids = [torch.randint(0, vocab, (n_ids,), generator=g, dtype=torch.int32) for _ in range(world)]
        for r in range(world):
            emu.ids[r][:n_ids].copy_(ids[r])
        emu.mark([emu.args(r, emu.mark_slot, 0, total, p2p.AR_TWO_SHOT, False, sst, n_ids=n_ids, dense_rows=dense_rows,
                           vocab_rows=vocab) for r in range(world)])
        if schedule == "one_launch":
            emu.allreduce([emu.args(r, 0, 0, total, p2p.AR_TWO_SHOT, True, sst, *state[r], rows=p2p.ROWS_IDLE)
                           for r in range(world)])
        # "backward": dense rows are written in full, word rows are scatter-added into the cleared live rows
        dense_g = torch.zeros(world, rows, d, device=dev)
        for r in range(world):
            gr = emu.g16[r].view(rows, d)
            torch.manual_seed(900 + 10 * step + r)
            dense_g[r, :dense_rows] = (torch.randn(dense_rows, d, device=dev) * 0.1).bfloat16().float()
            gr[:dense_rows] = dense_g[r, :dense_rows].bfloat16()
            upd = (torch.randn(n_ids, d, device=dev) * 0.1).bfloat16()
            for k, i in enumerate(ids[r].tolist()):  # same accumulation order as the reference below
                gr[dense_rows + i] += upd[k]
            live = torch.unique(torch.cat(ids).long().to(dev)) + dense_rows
            assert torch.isfinite(gr[live].float()).all(), "a live row was not cleared by the marking kernel"
            dense_g[r, live] = gr[live].float()
        if schedule == "one_launch":
            emu.allreduce([emu.args(r, 0, 0, total, p2p.AR_TWO_SHOT, True, sst, *state[r], rows=p2p.ROWS_LIVE)
                           for r in range(world)])
        else:
            emu.allreduce([emu.args(r, 0, 0, total, p2p.AR_TWO_SHOT, True, sst, *state[r], rows=p2p.ROWS_LIVE, phase=1)
                           for r in range(world)])
            emu.allreduce([emu.args(r, 0, 0, total, p2p.AR_TWO_SHOT, True, sst, *state[r], rows=p2p.ROWS_BOTH, phase=2,
                                    end_barrier=True) for r in range(world)])
        torch.cuda.synchronize()
        emu.check_errors()
        _ref_adam(K.read_step_state(sst), (dense_g.sum(0).view(-1) / world).bfloat16().float(), rm, rmm, rv)
        assert all(torch.equal(emu.stamps[0], s) for s in emu.stamps)
    assert all(torch.equal(emu.w16[0], w) for w in emu.w16)
    for r in range(world):
        assert torch.allclose(state[r][0], rm[r * n:(r + 1) * n], atol=2e-3, rtol=2e-3), r
        assert torch.allclose(state[r][1], rmm[r * n:(r + 1) * n], atol=1e-4, rtol=1e-2), r
        assert torch.allclose(state[r][2], rv[r * n:(r + 1) * n], atol=1e-6, rtol=2e-2), r
    assert (emu.w16[0].float() - rm).abs().max().item() <= 2 ** -8 * rm.abs().max().item() + 2e-3
