"""Multi-GPU worker run under torchrun by tests/test_multigpu.py (one process per GPU).
Each check prints 'OK <name>' or raises; rank 0's stdout is what the test asserts on."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import huggingface_sagemaker_tensorflow_distributed_b200.parallel.hvd as hvd  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.ops import kernels as K  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.parallel import comm, p2p  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.parallel.dp_engine import Bucket  # noqa: E402


def log(*a):
    if hvd.rank() == 0:
        print(*a, flush=True)


def gather_all(t):
    parts = [torch.empty_like(t) for _ in range(hvd.size())]
    dist.all_gather(parts, t.contiguous())
    return parts


def check_collectives():
    W, R, dev = hvd.size(), hvd.rank(), comm.device()
    sizes = [64 * W, 1024 * 8, 1 << 20, (3 << 20) + 128 * W]
    sizes = [s // (8 * W) * (8 * W) for s in sizes]
    starts, off = [], 0
    for s in sizes:
        starts.append(off)
        off += s
    buckets = [Bucket(i, st, st + n, []) for i, (st, n) in enumerate(zip(starts, sizes))]
    grp = p2p.P2PGroup(off, buckets, dev)
    log("arena", grp.arena.kind, "nvls", bool(grp.arena.mc_ptr))
    modes = [p2p.AR_TWO_SHOT, p2p.AR_ONE_SHOT] + ([p2p.AR_NVLS] if grp.arena.mc_ptr else [])
    for mode in modes:
        for rep in range(3):  # repeated calls exercise the epoch protocol
            torch.manual_seed(100 * rep + R)
            grp.g16.copy_((torch.randn(off, device=dev) * (1 + R)).bfloat16())
            grp.w16.zero_()
            torch.cuda.synchronize()
            dist.barrier()
            want = torch.stack([p.float() for p in gather_all(grp.g16)]).sum(0) / W
            dist.barrier()
            for b in buckets:
                grp.launch(grp.make_args(b.index, b.start, b.numel, mode, update=False))
            torch.cuda.synchronize()
            dist.barrier()
            got = grp.w16 if mode == p2p.AR_ONE_SHOT else grp.g16
            err = (got.float() - want).abs().max().item()
            assert err <= 0.02 * want.abs().max().item(), (p2p.MODE_NAMES[mode], rep, err)
            same = gather_all(got)
            assert all(torch.equal(same[0], s) for s in same), f"ranks disagree in {p2p.MODE_NAMES[mode]}"
            grp.check_errors()
        log("OK allreduce", p2p.MODE_NAMES[mode])
    # fused update vs. (reference all-reduce + flat AdamW)
    for mode in modes:
        grp2 = p2p.P2PGroup(off, buckets, dev, force_mode=mode)
        torch.manual_seed(7)
        w0 = torch.randn(off, device=dev)
        master = grp2.take_owned(w0)
        m, v = torch.zeros_like(master), torch.zeros_like(master)
        grp2.w16.copy_(w0.bfloat16())
        rm, rmm, rv = w0.clone(), torch.zeros(off, device=dev), torch.zeros(off, device=dev)
        rw = torch.empty(off, dtype=torch.bfloat16, device=dev)
        st = K.new_step_state(dev, seed=1)
        for step in range(3):
            torch.manual_seed(1000 + 10 * step + R)
            grp2.g16.copy_((torch.randn(off, device=dev) * 0.1).bfloat16())
            torch.cuda.synchronize()
            dist.barrier()
            avg = (torch.stack([p.float() for p in gather_all(grp2.g16)]).sum(0) / W)
            K.step_prologue(st, 1e-2, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.01)
            for b in buckets:
                _, _, o = grp2.owned[b.index]
                grp2.launch(grp2.make_args(b.index, b.start, b.numel, grp2.modes[b.index], True, st, master, m, v, o))
            torch.cuda.synchronize()
            dist.barrier()
            # the kernel rounds the average to fp32 (two-shot/one-shot) or bf16 (NVLS switch output)
            # (both sharded forms hand the average from their reduce pass to their optimizer pass in bf16)
            if mode == p2p.AR_ONE_SHOT:
                _ref_adam(st, avg, rm, rmm, rv, rw)
            else:
                K.adamw_flat(st, avg.bfloat16(), rm, rmm, rv, rw)
            torch.cuda.synchronize()
        full = grp2.gather_owned(master)
        tol = 1e-5 if mode == p2p.AR_ONE_SHOT else 2e-3
        assert torch.allclose(full, rm, atol=tol, rtol=tol), (p2p.MODE_NAMES[mode], (full - rm).abs().max().item())
        same = gather_all(grp2.w16)
        assert all(torch.equal(same[0], s) for s in same)
        assert (grp2.w16.float() - rm).abs().max().item() <= 0.02 * rm.abs().max().item()
        grp2.check_errors()
        log("OK fused_update", p2p.MODE_NAMES[mode])


def check_row_sparse(split=False):
    """one launch per part: mark -> idle rows -> live rows; split: mark -> reduce of live rows -> update of all rows.
    On real peers, either == dense reduce + AdamW (gathered fp32 reference)."""
    W, R, dev = hvd.size(), hvd.rank(), comm.device()
    d, dense_rows, vocab, n_ids = 256, 6, 2000, 160
    rows = dense_rows + vocab
    while (rows * d) % (8 * W):
        rows += 1
    vocab = rows - dense_rows
    total = rows * d
    b = Bucket(0, 0, total, [], sparse=True, row_elems=d, dense_rows=dense_rows, vocab_rows=vocab)
    grp = p2p.P2PGroup(total, [b], dev)
    mode = grp.modes[0]
    torch.manual_seed(3)
    w0 = torch.randn(total, device=dev)
    n = total // W
    master, m, v = w0[R * n:(R + 1) * n].clone(), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    grp.w16.copy_(w0.bfloat16())
    grp.g16.fill_(float("nan"))
    rm, rmm, rv = w0.clone(), torch.zeros(total, device=dev), torch.zeros(total, device=dev)
    rw = torch.empty(total, dtype=torch.bfloat16, device=dev)
    st = K.new_step_state(dev, seed=1)
    ids_slot = grp.ids_view(n_ids)
    for step in range(4):
        K.step_prologue(st, 1e-2, "constant", 0, 0, 0.9, 0.999, 1e-7, 0.0)
        g = torch.Generator(device="cpu").manual_seed(50 + step)
        ids = [torch.randint(0, vocab, (n_ids,), generator=g, dtype=torch.int32) for _ in range(W)]
        ids_slot.copy_(ids[R])
        mark = grp.make_args(grp.mark_slot, 0, total, mode, False, st, blocks=8)
        mark.n_ids = n_ids
        grp.launch_mark(mark)
        if not split:
            grp.launch(grp.make_args(0, 0, total, mode, True, st, master, m, v, 0, rows=p2p.ROWS_IDLE))
        gr = grp.g16.view(rows, d)
        torch.manual_seed(900 + 10 * step + R)
        gr[:dense_rows] = (torch.randn(dense_rows, d, device=dev) * 0.1).bfloat16()
        upd = (torch.randn(n_ids, d, device=dev) * 0.1).bfloat16()
        for k, i in enumerate(ids[R].tolist()):
            gr[dense_rows + i] += upd[k]
        live = torch.unique(torch.cat(ids).long().to(dev)) + dense_rows
        assert torch.isfinite(gr[live].float()).all(), "a live row was not cleared by the marking kernel"
        mine = torch.zeros(rows, d, device=dev)
        mine[:dense_rows] = gr[:dense_rows].float()
        mine[live] = gr[live].float()
        torch.cuda.synchronize()
        dist.barrier()
        avg = torch.stack(gather_all(mine)).sum(0).view(-1) / W
        if split:
            grp.launch(grp.make_args(0, 0, total, mode, True, st, master, m, v, 0, rows=p2p.ROWS_LIVE,
                                     blocks=grp.tail_blocks, tail=True, phase=1))
            grp.launch(grp.make_args(0, 0, total, mode, True, st, master, m, v, 0, rows=p2p.ROWS_BOTH,
                                     blocks=grp.tail_blocks, tail=True, phase=2, end_barrier=True))
        else:
            grp.launch(grp.make_args(0, 0, total, mode, True, st, master, m, v, 0, rows=p2p.ROWS_LIVE,
                                     blocks=grp.tail_blocks, tail=True))
        torch.cuda.synchronize()
        dist.barrier()
        K.adamw_flat(st, avg.bfloat16(), rm, rmm, rv, rw)  # the reduce pass hands a bf16 average to the optimizer pass
        torch.cuda.synchronize()
    grp.check_errors()
    tol = 2e-3
    full = grp.gather_owned(master)
    assert torch.allclose(full, rm, atol=tol, rtol=tol), (full - rm).abs().max().item()
    same = gather_all(grp.w16)
    assert all(torch.equal(same[0], s) for s in same), "weights differ across ranks"
    assert (grp.w16.float() - rm).abs().max().item() <= 0.02 * rm.abs().max().item()
    log("OK row_sparse" + (" split" if split else ""), p2p.MODE_NAMES[mode])


def _tiny_fit(samples, epochs=1, seed=5, on_step=None):
    from huggingface_sagemaker_tensorflow_distributed_b200 import trainer as T
    from huggingface_sagemaker_tensorflow_distributed_b200.data import BatchedDataset, synthetic_imdb
    from huggingface_sagemaker_tensorflow_distributed_b200.optim import Adam
    W, R, dev = hvd.size(), hvd.rank(), comm.device()
    model = T.AutoModelForSequenceClassification.from_pretrained("tiny-distilbert-gpu", device=dev, seed=seed, bucket_mb=0.5)
    ds = synthetic_imdb(model.cfg, samples, 128, seed=2)
    opt = hvd.DistributedOptimizer(Adam(learning_rate=1e-4 * W))
    model.compile(optimizer=opt, loss=T.SparseCategoricalCrossentropy(from_logits=True))
    bds = BatchedDataset(ds, 4, R, W)
    hist = model.fit(bds, epochs=epochs, verbose=0, log_every=2,
                     callbacks=[hvd.callbacks.BroadcastGlobalVariablesCallback(0)] + ([on_step] if on_step else []))
    return model, hist, len(bds)


def check_uneven():
    """65 samples: the strided shards differ by one sample; both ranks must run the same number of steps."""
    model, hist, steps = _tiny_fit(65)
    n = torch.tensor([steps], device=comm.device())
    parts = gather_all(n)
    assert all(int(p.item()) == steps for p in parts), [int(p.item()) for p in parts]
    loss = hist.history["loss"][0]
    assert loss == loss
    log("OK uneven shards", steps, "steps on every rank")


def check_dead_peer():
    """Rank 1 dies without warning after a few steps; rank 0 must fail with the watchdog error."""
    import signal
    import time
    from huggingface_sagemaker_tensorflow_distributed_b200 import trainer as T
    from huggingface_sagemaker_tensorflow_distributed_b200.data import BatchedDataset, synthetic_imdb
    from huggingface_sagemaker_tensorflow_distributed_b200.optim import Adam
    W, R, dev = hvd.size(), hvd.rank(), comm.device()
    model = T.AutoModelForSequenceClassification.from_pretrained("tiny-distilbert-gpu", device=dev, seed=5, bucket_mb=0.5)
    ds = synthetic_imdb(model.cfg, 4 * 4000 * W, 128, seed=2)
    opt = hvd.DistributedOptimizer(Adam(learning_rate=1e-4 * W))
    model.compile(optimizer=opt, loss=T.SparseCategoricalCrossentropy(from_logits=True))
    model.backend.setup_training(opt, 10 ** 6, 4, 128)
    model.broadcast_state(0)
    batches = list(BatchedDataset(ds, 4, R, W, max_batches=8))
    for i in range(8):
        model.backend.train_step(*batches[i])
    torch.cuda.synchronize()
    if R == 1:
        os.kill(os.getpid(), signal.SIGKILL)
    t0 = time.time()
    try:
        for i in range(100000):
            model.backend.train_step(*batches[i % 8])
            if (i + 1) % 20 == 0:
                model.backend.poll_health()
            if time.time() - t0 > 60:
                break
        model.backend.finish()
    except RuntimeError as e:
        print(f"rank {R}: {e}", flush=True)
        os._exit(3)
    print("rank 0 never noticed the dead peer", flush=True)
    os._exit(0)


def _ref_adam(st, g_fp32, w, m, v, wb):
    s = K.read_step_state(st)
    m.mul_(s.beta1).add_(g_fp32, alpha=1 - s.beta1)
    v.mul_(s.beta2).addcmul_(g_fp32, g_fp32, value=1 - s.beta2)
    w.mul_(s.decay).addcdiv_(m, v.sqrt() + s.eps, value=-s.lr_t)
    wb.copy_(w.bfloat16())


def check_training():
    from huggingface_sagemaker_tensorflow_distributed_b200 import trainer as T
    from huggingface_sagemaker_tensorflow_distributed_b200.data import BatchedDataset, synthetic_imdb
    from huggingface_sagemaker_tensorflow_distributed_b200.optim import Adam
    W, R, dev = hvd.size(), hvd.rank(), comm.device()
    losses = {}
    for comm_mode in ("p2p", "nccl"):
        model = T.AutoModelForSequenceClassification.from_pretrained(
            "tiny-distilbert-gpu", device=dev, dropout=0.0, seed=5 + R, comm_mode=comm_mode, bucket_mb=0.5)
        ds = synthetic_imdb(model.cfg, 16 * 4 * W, 128, seed=2)
        opt = hvd.DistributedOptimizer(Adam(learning_rate=1e-4 * W))
        model.compile(optimizer=opt, loss=T.SparseCategoricalCrossentropy(from_logits=True))
        hist = model.fit(BatchedDataset(ds, 4, R, W), epochs=2, verbose=0,
                         callbacks=[hvd.callbacks.BroadcastGlobalVariablesCallback(0)])
        losses[comm_mode] = hist.history["loss"]
        sd = model.state_dict()
        flat = torch.cat([v.reshape(-1) for v in sd.values()]).to(dev)
        same = gather_all(flat)
        assert all(torch.equal(same[0], s) for s in same), f"weights diverged across ranks ({comm_mode})"
        if comm_mode == "p2p":
            log("dp:", model.backend.dp.describe())
            assert model.backend.plans[(4, 128, True)].graph is not None
            p2p_flat = flat.clone()
        else:
            # Elements whose gradient is rounding noise move by +-lr per step under Adam in either run, so the
            # max-abs difference can approach the total movement (32 steps x 2e-4); the bulk must agree closely.
            diff = flat - p2p_flat
            travel = 32 * 1e-4 * W  # steps x learning rate (lr scales with the world size, scripts/train.py:112)
            assert diff.abs().max().item() < 2.1 * travel, diff.abs().max().item()
            assert (diff.norm() / flat.norm()).item() < 1e-2 * W, (diff.norm() / flat.norm()).item()
    lp, ln = losses["p2p"], losses["nccl"]
    assert abs(lp[-1] - ln[-1]) < 1e-2 * W, losses
    log("OK training p2p==nccl", lp, ln)


if __name__ == "__main__":
    hvd.init()
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "collectives"):
        check_collectives()
        check_row_sparse()
        check_row_sparse(split=True)
    if which == "uneven":
        check_uneven()
    if which == "dead_peer":
        check_dead_peer()
    if which in ("all", "training"):
        check_training()
    dist.barrier()
    log("ALL OK")
    hvd.shutdown()
