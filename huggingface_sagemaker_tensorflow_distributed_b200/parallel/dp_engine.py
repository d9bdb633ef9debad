"""Data-parallel gradient engine: bucket plan, arenas, optimizer placement
(reference scripts/train.py:114,133).

Horovod's C++ core negotiates ready tensors, packs them into a fusion buffer and calls
``ncclAllReduce`` + postscale; Keras then applies Adam per variable.  Here the gradient arena
*is* the fusion buffer (parameters are laid out in backward-ready order, so a bucket is a
contiguous byte range that completes front-to-back), and reduce + 1/world scale +
bf16→fp32 cast + Keras-AdamW + bf16 weight write-back are one kernel per bucket:

  * world_size == 1 : the flat AdamW kernel over the whole arena (no communication);
  * mode "p2p"      : hand-written NVLink kernels over IPC-mapped peer arenas
                      (csrc/allreduce_sm100.cu) — one-shot / two-shot / NVLS by bucket
                      size, launched on a side stream as each bucket's gradients complete;
  * mode "nccl"     : library all-reduce + the flat AdamW kernel — the baseline plumbing,
                      kept for A/B measurements.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Tuple

import torch
import torch.distributed as dist

from ..ops import kernels as K
from . import comm


@dataclass
class Bucket:
    index: int
    start: int  # element offsets into the arenas
    end: int
    names: List[str]
    # row-sparse bucket (the embedding tables): a stack of rows of `row_elems` elements, `dense_rows` always-live
    # rows (LayerNorm / token-type / position tables) followed by `vocab_rows` word-embedding rows
    sparse: bool = False
    row_elems: int = 0
    dense_rows: int = 0
    vocab_rows: int = 0

    @property
    def numel(self) -> int:
        return self.end - self.start


P2P_WORLD_SIZES = (2, 4, 8)
HEAD_PARAMS = ("head.out.weight", "head.out.bias", "head.dense.weight", "head.dense.bias")


def plan_buckets(slots: Dict[str, "object"], order: List[str], bucket_elems: int, world: int) -> List[Bucket]:
    """Greedy contiguous bucketing in backward-ready order, done by the native bucket manager
    (csrc/bucket_manager.cpp: hstd_bucket_plan).  Bucket boundaries fall on parameter
    boundaries; sizes are padded by the arena alignment so every bucket splits into ``world``
    16-byte-aligned slices."""
    from ..ops import host

    spans, param_bucket = host.bucket_plan([slots[n].offset for n in order], [slots[n].end for n in order],
                                           bucket_elems)
    buckets = [Bucket(i, a, b, []) for i, (a, b) in enumerate(spans)]
    for n, bi in zip(order, param_bucket):
        buckets[bi].names.append(n)
    return buckets


def plan_buckets_with_embeddings(slots, order: List[str], bucket_elems: int, world: int, row_sparse: bool) -> List[Bucket]:
    """Encoder + head parameters are bucketed greedily; the embedding block (LayerNorm, token-type, position and
    word tables — the tail of the readiness order, complete only when backward ends) is one bucket of its
    own.  With ``row_sparse`` it is reduced row-sparsely: all its tensors are stacks of hidden-size rows."""
    emb = [n for n in order if n.startswith("emb.")]
    rest = [n for n in order if not n.startswith("emb.")]
    assert order[-len(emb):] == emb and emb[-1] == "emb.word.weight"
    # The gradients that complete last among the encoder's (layer 0: attention output + QKV projections) get a
    # small bucket of their own: whatever is reduced after them is no longer hidden by compute, so the tail
    # should not also carry the leftovers of the greedy packing.
    last = [n for n in rest if n.startswith("layers.0.") and n.split(".")[2] in ("ln1", "out", "qkv")]
    if last and rest[-len(last):] == last and len(rest) > len(last):
        buckets = plan_buckets(slots, rest[:-len(last)], bucket_elems, world)
        buckets.append(Bucket(len(buckets), slots[last[0]].offset, slots[last[-1]].end, last))
    else:
        buckets = plan_buckets(slots, rest, bucket_elems, world)
    word = slots["emb.word.weight"]
    d = word.shape[1]
    start, end = slots[emb[0]].offset, word.end
    b = Bucket(len(buckets), start, end, emb)
    contiguous_rows = all(slots[n].numel % d == 0 and slots[n].end - slots[n].offset == slots[n].numel for n in emb)
    if row_sparse and contiguous_rows and (end - start) % (8 * world) == 0:
        b.sparse, b.row_elems = True, d
        b.dense_rows, b.vocab_rows = (word.offset - start) // d, word.shape[0]
    buckets.append(b)
    return buckets


class DPEngine:
    def __init__(self, engine, bucket_mb: float = 32.0, mode: str = "p2p"):
        self.engine = engine
        self.device = engine.device
        self.world, self.rank = comm.size(), comm.rank()
        self.mode = mode if self.world > 1 else "local"
        if self.mode == "p2p" and self.world not in P2P_WORLD_SIZES:
            # the fused kernels are instantiated for 2, 4 and 8 ranks (csrc/allreduce_sm100.cu); a box with
            # 3, 5, 6 or 7 visible GPUs still trains, on the library all-reduce + flat AdamW path
            import warnings

            warnings.warn(f"fused NVLink all-reduce supports world sizes {P2P_WORLD_SIZES}, got {self.world}: "
                          "falling back to comm mode 'nccl'")
            self.mode = "nccl"
        self.total = engine.total
        order = list(engine.slots.keys())
        import os

        row_sparse = self.mode == "p2p" and os.environ.get("HSTD_SPARSE_EMB", "1") != "0"
        self.buckets = plan_buckets_with_embeddings(engine.slots, order, int(bucket_mb * (1 << 20) / 2), self.world,
                                                    row_sparse)
        self.name_to_bucket = {n: b.index for b in self.buckets for n in b.names}
        from ..ops import host

        # readiness bookkeeping lives in the native tracker; head gradients are produced first and
        # never announced separately, so they are exempt
        self._param_index = {n: i for i, n in enumerate(order)}
        self._tracker = host.BucketTracker([self.name_to_bucket[n] for n in order],
                                           [n in HEAD_PARAMS for n in order], len(self.buckets))
        self.distributed = False
        self._pending_full = None
        self._opt_stream = None
        self._alloc()

    # ------------------------------------------------------------------ memory
    def _alloc(self) -> None:
        dev = self.device
        self.p2p = None
        if self.mode == "p2p":
            from .p2p import P2PGroup

            self.p2p = P2PGroup(self.total, self.buckets, dev)
            self.w16, self.g16 = self.p2p.w16, self.p2p.g16
        else:
            self.w16 = torch.zeros(self.total, dtype=torch.bfloat16, device=dev)
            self.g16 = torch.zeros(self.total, dtype=torch.bfloat16, device=dev)
        # fp32 master + Adam moments.  In p2p mode each rank keeps only the slice of every
        # bucket it owns (ZeRO-1-like); stored compactly in bucket order.
        if self.mode == "p2p":
            n = self.p2p.owned_elems
        else:
            n = self.total
        self.master = torch.zeros(n, dtype=torch.float32, device=dev)
        self.m = torch.zeros(n, dtype=torch.float32, device=dev)
        self.v = torch.zeros(n, dtype=torch.float32, device=dev)

    def configure(self, distributed: bool) -> None:
        self.distributed = distributed and self.world > 1

    def wants_early_buckets(self) -> bool:
        """True when a bucket is consumed as soon as its gradients exist: the fused all-reduce
        (distributed p2p) or, on a single GPU, the per-bucket AdamW on the optimizer stream."""
        return (self.p2p is not None and self.distributed) or self._local_overlap()

    def local_early_buckets(self) -> bool:
        return self._local_overlap()

    def _local_overlap(self) -> bool:
        # Measured on B200 (DistilBERT, 1 GPU): overlapping the per-bucket AdamW with backward is a
        # net loss (2.63 vs 2.56 ms/step) — its HBM traffic slows the concurrent GEMMs more than it
        # hides — so the single-GPU default stays "one flat AdamW at the end of the step".
        import os
        return (self.mode == "local" and not self.distributed and os.environ.get("HSTD_LOCAL_OPT_OVERLAP") == "1")

    def capturable(self) -> bool:
        return not (self.distributed and self.mode == "nccl")

    def check_errors(self) -> None:
        if self.p2p is not None:
            self.p2p.check_errors()

    def poll_errors(self) -> None:
        if self.p2p is not None:
            self.p2p.poll_errors()

    def describe(self) -> dict:
        d = {"mode": self.mode, "world": self.world, "num_buckets": len(self.buckets)}
        if self.p2p is not None:
            d.update(self.p2p.describe())
        return d

    # ------------------------------------------------------------------ state movement
    def set_master(self, flat_fp32: torch.Tensor) -> None:
        """Install full fp32 weights.  With sharded optimizer state a rank keeps fp32 only for
        the slices it owns, so until training starts a host copy of what *this* rank was given
        is retained: a later broadcast_state(root) can then ship root's complete fp32 tensor
        even if ranks were initialised differently."""
        self._pending_full = flat_fp32.detach().cpu() if self.mode == "p2p" else None
        self.w16.copy_(flat_fp32.to(torch.bfloat16))
        if self.mode == "p2p":
            self.master.copy_(self.p2p.take_owned(flat_fp32))
        else:
            self.master.copy_(flat_fp32)
        torch.cuda.synchronize()
        comm.barrier()

    def gather_master(self) -> torch.Tensor:
        if self.mode == "p2p":
            return self.p2p.gather_owned(self.master)
        return self.master.clone()

    def gather_moments(self) -> Tuple[torch.Tensor, torch.Tensor]:
        if self.mode == "p2p":
            return self.p2p.gather_owned(self.m), self.p2p.gather_owned(self.v)
        return self.m.clone(), self.v.clone()

    def set_moments(self, m: torch.Tensor, v: torch.Tensor) -> None:
        if self.mode == "p2p":
            self.m.copy_(self.p2p.take_owned(m))
            self.v.copy_(self.p2p.take_owned(v))
        else:
            self.m.copy_(m)
            self.v.copy_(v)

    def broadcast_state(self, root: int, step_state: torch.Tensor) -> None:
        """Rank-``root`` state to everyone (BroadcastGlobalVariablesCallback,
        scripts/train.py:127-134).  One-time, so NCCL is fine here (K12)."""
        if self.world <= 1:
            return
        pending = getattr(self, "_pending_full", None)
        full = pending.to(self.device) if pending is not None else self.gather_master()
        dist.broadcast(full, src=root)
        mm, vv = self.gather_moments()
        dist.broadcast(mm, src=root)
        dist.broadcast(vv, src=root)
        self.set_master(full)
        self.set_moments(mm, vv)
        dist.broadcast(step_state, src=root)

    # ------------------------------------------------------------------ hooks used while recording a plan
    def sparse_embedding(self) -> bool:
        """True when the embedding bucket is reduced row-sparsely (distributed p2p training)."""
        return self.p2p is not None and self.distributed and self.p2p.sparse is not None

    def ids_buffer(self, n: int):
        """Peer-visible slot for the step's token ids (row-sparse embedding reduce), else None."""
        return self.p2p.ids_view(n) if self.sparse_embedding() else None

    def after_step_prologue(self, plan, n_ids: int) -> None:
        if self.sparse_embedding():
            self.p2p.record_step_start(plan, self.engine.state, self.master, self.m, self.v, n_ids)

    def before_embedding_scatter(self, plan) -> None:
        if self.sparse_embedding():
            self.p2p.record_wait_rows_marked(plan)

    def begin_backward(self, plan) -> None:
        self._tracker.reset()
        if self.p2p is not None and self.distributed:
            self.p2p.record_begin(plan)

    def after_grads(self, plan, names: List[str]) -> None:
        """Called (at record time) once the gradients of ``names`` have been produced.  The
        tracker releases buckets strictly in order because the arena is in readiness order."""
        ready = [self.buckets[i] for i in self._tracker.mark([self._param_index[n] for n in names])]
        if self.p2p is not None and self.distributed:
            side = self.engine.side_stream if getattr(plan, "side_used", False) else None
            for b in ready:
                self.p2p.record_bucket(plan, b, self.engine.state, self.master, self.m, self.v,
                                       overlap=self.engine.overlap, side_stream=side,
                                       last=b.index >= len(self.buckets) - 2)
        elif self._local_overlap():
            # single GPU: the HBM-bound AdamW of a finished bucket runs on its own stream under the
            # tensor-bound rest of backward instead of serially at the end of the step
            for b in ready:
                self._record_local_bucket(plan, b)

    def _record_local_bucket(self, plan, b) -> None:
        if self._opt_stream is None:
            self._opt_stream = torch.cuda.Stream(device=self.device)
        os_, st = self._opt_stream, self.engine.state
        ev = torch.cuda.Event()
        plan.keep.append(ev)
        sl = slice(b.start, b.end)
        g, w, mm, vv, wb = self.g16[sl], self.master[sl], self.m[sl], self.v[sl], self.w16[sl]

        def op():
            ev.record(torch.cuda.current_stream())
            os_.wait_event(ev)
            K.adamw_flat(st, g, w, mm, vv, wb, 1.0, stream=os_)

        plan.launches += 1
        plan.ops.append(op)
        plan.names.append("adamw(bucket)@opt")
        plan.opt_dirty = True

    def end_backward(self, plan) -> None:
        self._pending_full = None  # training is about to start: sharded state is now the truth
        assert self._tracker.done == len(self.buckets), "every bucket must be complete at end of backward"
        st = self.engine.state
        if self.p2p is not None and self.distributed:
            self.p2p.record_end(plan, st, self.master, self.m, self.v, overlap=self.engine.overlap)
            return
        if self.mode == "p2p":
            raise NotImplementedError(
                "world_size > 1 with the p2p arenas shards the optimizer state across ranks, so the optimizer must "
                "be wrapped in hvd.DistributedOptimizer (or pass comm='nccl' for independent replicas)")
        if self._local_overlap():
            if getattr(plan, "opt_dirty", False):
                os_ = self._opt_stream
                ev = torch.cuda.Event()
                plan.keep.append(ev)

                def join():
                    ev.record(os_)
                    torch.cuda.current_stream().wait_event(ev)

                plan.ops.append(join)
                plan.names.append("join(opt)")
            return
        if self.distributed:  # nccl baseline plumbing
            g = self.g16
            plan.launches += 1
            plan.ops.append(lambda: dist.all_reduce(g))
            plan.names.append("nccl all_reduce")
            scale = 1.0 / self.world
        else:
            scale = 1.0
        plan.launches += 1
        plan.ops.append(lambda: K.adamw_flat(st, self.g16, self.master, self.m, self.v, self.w16, scale))
        plan.names.append("adamw(flat)")
