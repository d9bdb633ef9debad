"""B200 execution engine: the encoder classifier as a hand-scheduled forward/backward over
the sm_100a kernel library, with flat parameter arenas and CUDA-graph replay.

Design (B200-first, not autograd):
  * **Flat arenas.**  All parameters live in one bf16 weight arena and one bf16 gradient
    arena, ordered by *backward readiness* (head, layer L-1 … layer 0, embeddings) so the
    gradient buckets the data-parallel engine reduces are contiguous ranges that become
    ready front-to-back during backward.  fp32 master weights and Adam moments mirror the
    layout (sharded per bucket slice across ranks when world_size > 1).
  * **Static plans.**  For each (batch, seq, train) shape the engine pre-allocates every
    activation/scratch buffer and records the step as a flat list of kernel calls with
    pre-marshalled argument structs; the list is captured once into a CUDA graph and
    replayed — no allocator, autograd or Python work on the hot path.
  * **Kernels.**  Every GEMM (QKV, attention QKᵀ / PV and their four backward products,
    out-proj, FFN, dgrad, wgrad) is the tcgen05/TMEM kernel; LayerNorm / softmax / GELU /
    dropout / loss are the fused row kernels; the optimizer is fused into the gradient
    reduction (parallel/dp_engine.py).

It implements the math the reference gets from TFAutoModelForSequenceClassification +
Keras fit/evaluate (scripts/train.py:117-123,145,170); the per-op mapping is SURVEY.md §2.4.
"""
from __future__ import annotations

import os

import math
from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Tuple

import torch

from ..ops import _lib, gemm as G, kernels as K
from ..optim import Adam
from ..parallel import comm
from .configuration import EncoderConfig
from .hf_names import canonical_names
from .reference import EncoderForSequenceClassification

ALIGN = 128  # elements; keeps every parameter 256-byte aligned inside the arenas
METRICS_OFFSET = 40  # byte offset of hstd_step_state.metrics
METRICS_ACC_OFFSET = 56


@dataclass
class ParamSlot:
    name: str
    shape: Tuple[int, ...]
    offset: int
    numel: int

    @property
    def end(self) -> int:
        return self.offset + ((self.numel + ALIGN - 1) // ALIGN) * ALIGN


def backward_ready_order(cfg: EncoderConfig) -> List[str]:
    """Canonical parameter names in the order their gradients complete during backward."""
    names = ["head.out.weight", "head.out.bias", "head.dense.weight", "head.dense.bias"]
    for i in reversed(range(cfg.num_layers)):
        p = f"layers.{i}."
        names += [p + "ln2.weight", p + "ln2.bias", p + "ffn2.weight", p + "ffn2.bias", p + "ffn1.weight",
                  p + "ffn1.bias", p + "ln1.weight", p + "ln1.bias", p + "out.weight", p + "out.bias",
                  p + "qkv.weight", p + "qkv.bias"]
    names += ["emb.ln.weight", "emb.ln.bias"]
    if cfg.type_vocab_size > 0:
        names.append("emb.tok_type.weight")
    names += ["emb.pos.weight", "emb.word.weight"]
    assert sorted(names) == sorted(canonical_names(cfg))
    return names


def param_shapes(cfg: EncoderConfig) -> Dict[str, Tuple[int, ...]]:
    with torch.device("meta"):
        m = EncoderForSequenceClassification(cfg)
    return {k: tuple(v.shape) for k, v in m.state_dict().items()}


class Plan:
    """Buffers + recorded op list (+ CUDA graph) for one (B, S, train) shape."""

    def __init__(self, B: int, S: int, train: bool):
        self.B, self.S, self.train = B, S, train
        self.ops: List[Callable[[], None]] = []
        self.names: List[str] = []  # one label per op (profiling / NVTX)
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.bufs: Dict[str, torch.Tensor] = {}
        self.keep: list = []  # events referenced by recorded ops
        self.launches = 0  # kernel launches per step (counted while recording)
        self.warm = 0


class EngineBackend:
    name = "engine"

    def __init__(self, cfg: EncoderConfig, device: torch.device, seed: int = 42, bucket_mb: float = 32.0,
                 comm_mode: str = "p2p", use_cuda_graph: bool = True, overlap: bool = True,
                 fused_attention: bool = True, dual_stream: bool = True):
        if device.type != "cuda":
            raise RuntimeError("EngineBackend needs a CUDA device")
        _lib.load()  # fail loudly if libhstd_b200.so is missing: no silent eager fallback
        if cfg.hidden_size % 256 or cfg.hidden_size > 1024:
            raise ValueError("engine kernels need hidden_size in {256, 512, 768, 1024}")
        if cfg.head_dim != 64:
            raise ValueError("engine attention path is built for head_dim == 64")
        self.cfg, self.device = cfg, device
        self.use_cuda_graph = use_cuda_graph
        self.bucket_mb, self.comm_mode, self.overlap = bucket_mb, comm_mode, overlap
        self.fused_attention = fused_attention
        # weight-gradient GEMMs run on a second stream, concurrently with the dgrad chain
        self.dual_stream = dual_stream
        self.side_stream = torch.cuda.Stream(device=device) if dual_stream else None
        self.seed = seed
        self.world, self.rank = comm.size(), comm.rank()

        shapes = param_shapes(cfg)
        self.slots: Dict[str, ParamSlot] = {}
        off = 0
        for n in backward_ready_order(cfg):
            numel = math.prod(shapes[n])
            s = ParamSlot(n, shapes[n], off, numel)
            self.slots[n] = s
            off = s.end
        self.total = off
        # the data-parallel engine owns arena allocation (symmetric / IPC-visible when world > 1)
        from ..parallel.dp_engine import DPEngine

        self.dp = DPEngine(self, bucket_mb=bucket_mb, mode=comm_mode)
        self.w16 = self.dp.w16
        self.g16 = self.dp.g16
        self.state = K.new_step_state(device, seed=seed)
        self._init_weights(seed)
        self.plans: Dict[Tuple[int, int, bool], Plan] = {}
        self.opt: Optional[Adam] = None
        self.distributed = False
        self._metrics_view = self.state[METRICS_ACC_OFFSET:METRICS_ACC_OFFSET + 32].view(torch.float64)
        self._scratch: Dict[str, torch.Tensor] = {}

    # ------------------------------------------------------------------ parameters
    def W(self, name: str) -> torch.Tensor:
        s = self.slots[name]
        return self.w16[s.offset:s.offset + s.numel].view(s.shape)

    def Gr(self, name: str) -> torch.Tensor:
        s = self.slots[name]
        return self.g16[s.offset:s.offset + s.numel].view(s.shape)

    def _init_weights(self, seed: int) -> None:
        """Same initialisation as the PyTorch oracle with the same seed."""
        rng = torch.random.get_rng_state()
        torch.manual_seed(seed)
        ref = EncoderForSequenceClassification(self.cfg)
        torch.random.set_rng_state(rng)
        self.load_state_dict(ref.state_dict())

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        flat = torch.zeros(self.total, dtype=torch.float32)
        cur = None
        for n, s in self.slots.items():
            if n in sd:
                flat[s.offset:s.offset + s.numel] = sd[n].detach().float().reshape(-1).cpu()
            else:
                if cur is None:
                    cur = self.dp.gather_master().cpu()
                flat[s.offset:s.offset + s.numel] = cur[s.offset:s.offset + s.numel]
        self.dp.set_master(flat.to(self.device))

    def state_dict(self) -> Dict[str, torch.Tensor]:
        flat = self.dp.gather_master().cpu()
        return {n: flat[s.offset:s.offset + s.numel].view(s.shape).clone() for n, s in self.slots.items()
                if n in canonical_names(self.cfg)}

    def optimizer_state(self) -> dict:
        st = K.read_step_state(self.state)
        m, v = self.dp.gather_moments()
        return {"iterations": int(st.step), "m": m.cpu(), "v": v.cpu()}

    def load_optimizer_state(self, st: dict) -> None:
        if not st or "m" not in st:
            return  # (a PyTorch-backend checkpoint keeps per-parameter moments; only weights carry over)
        self.dp.set_moments(st["m"].to(self.device), st["v"].to(self.device))
        self.state = K.new_step_state(self.device, seed=self.seed, step=int(st.get("iterations", 0)))
        self._metrics_view = self.state[METRICS_ACC_OFFSET:METRICS_ACC_OFFSET + 32].view(torch.float64)
        self.plans.clear()

    def broadcast_state(self, root: int = 0) -> None:
        self.dp.broadcast_state(root, self.state)

    # ------------------------------------------------------------------ training setup
    def setup_training(self, opt: Adam, total_steps: int, batch_size: int, seq_len: int) -> None:
        self.opt = opt
        self.sched = opt.resolve_schedule(total_steps)
        self.distributed = bool(opt.distributed) and self.world > 1
        self.dp.configure(distributed=self.distributed)
        for key in [k for k in self.plans if k[2]]:
            del self.plans[key]  # hyper-parameters are baked into recorded ops

    # ------------------------------------------------------------------ plan building
    def _buf(self, plan: Plan, name: str, shape, dtype=torch.bfloat16) -> torch.Tensor:
        t = torch.empty(shape, dtype=dtype, device=self.device)
        plan.bufs[name] = t
        return t

    def _shared(self, name: str, numel: int, dtype=torch.float32) -> torch.Tensor:
        t = self._scratch.get(name)
        if t is None or t.numel() < numel or t.dtype != dtype:
            t = torch.empty(numel, dtype=dtype, device=self.device)
            self._scratch[name] = t
        return t

    def _gemm_op(self, plan: Plan, **kw) -> None:
        """Record one GEMM.  kw: A,B,C (tensors or (tensor, elem_offset)), strides, dims..."""
        g = _lib.GemmArgs()

        def ptr(x):
            if x is None:
                return None
            if isinstance(x, tuple):
                return x[0].data_ptr() + 2 * x[1]
            return x.data_ptr()

        g.A, g.B, g.C, g.bias, g.aux = ptr(kw["A"]), ptr(kw["B"]), ptr(kw["C"]), ptr(kw.get("bias")), ptr(kw.get("aux"))
        g.lda, g.ldb, g.ldc, g.ldaux = kw["lda"], kw["ldb"], kw["ldc"], kw.get("ldaux", 0)
        g.a_bs0, g.a_bs1 = kw.get("a_bs0", 0), kw.get("a_bs1", 0)
        g.b_bs0, g.b_bs1 = kw.get("b_bs0", 0), kw.get("b_bs1", 0)
        g.c_bs0, g.c_bs1 = kw.get("c_bs0", 0), kw.get("c_bs1", 0)
        g.M, g.N, g.K = kw["M"], kw["N"], kw["K"]
        g.nb0, g.nb1 = kw.get("nb0", 1), kw.get("nb1", 1)
        g.a_mn_major, g.b_mn_major = int(kw.get("a_t", False)), int(kw.get("b_t", False))
        g.epilogue, g.alpha = kw.get("epilogue", G.EPI_NONE), kw.get("alpha", 1.0)
        g.block_n, g.max_ctas = kw.get("block_n", 0), kw.get("max_ctas", 0)
        splits = kw.get("splits", 1)
        if g.nb0 == 1 and g.nb1 == 1 and not kw.get("block_n"):
            tuned = G.tuned_config(g.M, g.N, g.K, bool(g.a_mn_major), bool(g.b_mn_major))
            if tuned is not None:
                g.block_n, cm, tsplits = tuned
                g.cluster_m = 1 if cm == 1 else 0
                if g.a_mn_major and g.b_mn_major and g.epilogue == G.EPI_NONE and g.bias is None:
                    splits = tsplits
        ws = None
        if splits > 1:
            ws = self._shared("splitk_ws", g.M * g.N * splits, torch.float32)
            plan.launches += 1
        plan.launches += 1
        plan.ops.append(lambda g=g, splits=splits, ws=ws: G.gemm_raw(g, None, splits, ws))
        plan.names.append(kw.get("name") or f"gemm[{g.M}x{g.N}x{g.K}{'T' if g.a_mn_major else 'N'}{'T' if g.b_mn_major else 'N'}"
                          f"{'/b%d' % (g.nb0 * g.nb1) if g.nb0 * g.nb1 > 1 else ''}]")

    def _linear_fwd(self, plan, x, w_name, out, M, epilogue=G.EPI_NONE, aux=None):
        w = self.W(w_name + ".weight")
        N, Kd = w.shape
        self._gemm_op(plan, A=x, B=w, C=out, bias=self.W(w_name + ".bias"), lda=Kd, ldb=Kd, ldc=N, M=M, N=N, K=Kd,
                      epilogue=epilogue, aux=aux, ldaux=N)

    def _on_side(self, plan, first: int) -> None:
        """Re-home the ops recorded since index ``first`` onto the side stream: they start once
        everything recorded on the main stream so far has finished (event fork)."""
        if not self.dual_stream:
            return
        side = self.side_stream
        ev = torch.cuda.Event()
        plan.keep.append(ev)
        moved = plan.ops[first:]
        del plan.ops[first:]
        names = plan.names[first:]
        del plan.names[first:]

        def fork(ev=ev):
            ev.record(torch.cuda.current_stream())
            side.wait_event(ev)

        plan.ops.append(fork)
        plan.names.append("fork")
        for f, n in zip(moved, names):
            def run(f=f):
                with torch.cuda.stream(side):
                    f()
            plan.ops.append(run)
            plan.names.append(n + "@side")
        plan.side_dirty = True
        plan.side_used = True  # the side stream is part of this plan's captured graph from here on

    def _join_side(self, plan) -> None:
        """Main stream waits for everything queued on the side stream so far."""
        if not self.dual_stream or not getattr(plan, "side_dirty", False):
            return
        side = self.side_stream
        ev = torch.cuda.Event()
        plan.keep.append(ev)

        def join(ev=ev):
            ev.record(side)
            torch.cuda.current_stream().wait_event(ev)

        plan.ops.append(join)
        plan.names.append("join")
        plan.side_dirty = False

    def _linear_bwd(self, plan, dy, x, w_name, dx, M, *, epilogue=G.EPI_NONE, aux=None, ldaux=0, bias_grad=True):
        """dW = dyᵀ·x ; db = colsum(dy) (unless fused elsewhere) ; dx = dy·W (+ epilogue).
        dW / db do not feed the rest of backward, so they go to the side stream."""
        w = self.W(w_name + ".weight")
        N, Kd = w.shape  # forward: [M,Kd] x [N,Kd]ᵀ
        tiles = ((N + 127) // 128) * ((Kd + 127) // 128)
        splits = max(1, min(8, round(148 / tiles), M // 512))
        first = len(plan.ops)
        self._gemm_op(plan, A=dy, B=x, C=self.Gr(w_name + ".weight"), lda=N, ldb=Kd, ldc=Kd, M=N, N=Kd, K=M, a_t=True,
                      b_t=True, splits=splits)
        if bias_grad:
            # one scratch per stream: main-stream colsums (token-type grads) must not share it
            part = self._shared("colsum_partials_side", K.colsum_partials_floats(N))
            db = self.Gr(w_name + ".bias")
            plan.launches += 2
            plan.ops.append(lambda dy=dy, db=db, part=part: K.colsum(dy, db, part))
            plan.names.append("colsum(bias grad)")
        self._on_side(plan, first)
        if dx is not None:
            self._gemm_op(plan, A=dy, B=w, C=dx, lda=N, ldb=Kd, ldc=Kd, M=M, N=Kd, K=N, b_t=True, epilogue=epilogue,
                          aux=aux, ldaux=ldaux)

    def _attention_bwd_unfused(self, plan, i, qkv, Pd, P, dP, dctx, dqkv, B, S, H, dh, d, scale, p_att, st, op):
        """Attention backward as four batched GEMMs + a softmax-backward kernel (materialised P)."""
        hb = dict(nb0=H, nb1=B)
        q_s = dict(bs0=dh, bs1=S * 3 * d)  # strides of a Q/K/V/dQ/dK/dV slice inside [M,3d]
        # dPd = dctx · Vᵀ
        self._gemm_op(plan, A=dctx, B=(qkv[i], 2 * d), C=dP, lda=d, ldb=3 * d, ldc=S, a_bs0=dh, a_bs1=S * d,
                      b_bs0=q_s["bs0"], b_bs1=q_s["bs1"], c_bs0=S * S, c_bs1=H * S * S, M=S, N=S, K=dh, **hb)
        # dV = Pdᵀ · dctx
        self._gemm_op(plan, A=Pd[i], B=dctx, C=(dqkv, 2 * d), lda=S, ldb=d, ldc=3 * d, a_bs0=S * S, a_bs1=H * S * S,
                      b_bs0=dh, b_bs1=S * d, c_bs0=q_s["bs0"], c_bs1=q_s["bs1"], M=S, N=dh, K=S, a_t=True, b_t=True,
                      **hb)
        op(lambda i=i: K.softmax_bwd(P[i], dP, p_drop=p_att, site=i * 8 + 1, state=st))
        # dQ = scale * dS · K
        self._gemm_op(plan, A=dP, B=(qkv[i], d), C=dqkv, lda=S, ldb=3 * d, ldc=3 * d, a_bs0=S * S, a_bs1=H * S * S,
                      b_bs0=q_s["bs0"], b_bs1=q_s["bs1"], c_bs0=q_s["bs0"], c_bs1=q_s["bs1"], M=S, N=dh, K=S,
                      b_t=True, alpha=scale, **hb)
        # dK = scale * dSᵀ · Q
        self._gemm_op(plan, A=dP, B=qkv[i], C=(dqkv, d), lda=S, ldb=3 * d, ldc=3 * d, a_bs0=S * S, a_bs1=H * S * S,
                      b_bs0=q_s["bs0"], b_bs1=q_s["bs1"], c_bs0=q_s["bs0"], c_bs1=q_s["bs1"], M=S, N=dh, K=S,
                      a_t=True, b_t=True, alpha=scale, **hb)

    def build_plan(self, B: int, S: int, train: bool) -> Plan:
        cfg = self.cfg
        if S % 128 or S > 512:
            raise ValueError("engine needs seq_len in {128, 256, 384, 512}")
        plan = Plan(B, S, train)
        d, F, H, dh, L = cfg.hidden_size, cfg.intermediate_size, cfg.num_heads, cfg.head_dim, cfg.num_layers
        M = B * S
        st = self.state
        p_hid = cfg.hidden_dropout if train else 0.0
        p_att = cfg.attention_dropout if train else 0.0
        p_attn_out = p_hid if cfg.model_type != "distilbert" else 0.0
        p_cls = cfg.classifier_dropout if train else 0.0
        scale = 1.0 / math.sqrt(dh)
        i32 = torch.int32
        ids = self._buf(plan, "ids", (B, S), i32)
        if train:
            shared_ids = self.dp.ids_buffer(B * S)  # row-sparse embedding reduce: peers read this step's ids
            if shared_ids is not None:
                ids = plan.bufs["ids"] = shared_ids.view(B, S)
        mask = self._buf(plan, "mask", (B, S), i32)
        labels = self._buf(plan, "labels", (B,), i32)
        pos = self._buf(plan, "pos", (B, S), i32)
        nl = L if train else 1  # eval keeps one layer's worth of buffers

        def lb(name, shape, dtype=torch.bfloat16):
            return [self._buf(plan, f"{name}.{i}", shape, dtype) for i in range(nl)]

        x_in = lb("x", (M, d)) + [self._buf(plan, "x_final", (M, d))] if train else None
        if not train:
            xa, xb = self._buf(plan, "xa", (M, d)), self._buf(plan, "xb", (M, d))
        qkv, ctx = lb("qkv", (M, 3 * d)), lb("ctx", (M, d))
        fused = self.fused_attention
        if fused:
            P = Pd = None
            lse = lb("lse", (B, H, S), torch.float32)
        else:
            P = lb("P", (B, H, S, S))
            Pd = lb("Pd", (B, H, S, S)) if p_att > 0 else P
        r1, h1, r2 = lb("r1", (M, d)), lb("h1", (M, d)), lb("r2", (M, d))
        u, f = lb("u", (M, F)), lb("f", (M, F))
        f32 = torch.float32
        mean1, rstd1, mean2, rstd2 = (lb(n, (M,), f32) for n in ("mean1", "rstd1", "mean2", "rstd2"))
        pre0 = self._buf(plan, "pre0", (M, d))
        mean0, rstd0 = self._buf(plan, "mean0", (M,), f32), self._buf(plan, "rstd0", (M,), f32)
        eps = cfg.layer_norm_eps
        sref = st if train else None

        def op(fn, n_launch=1, name=None):
            plan.launches += n_launch
            plan.ops.append(fn)
            if name is None:  # label = first kernel-wrapper name referenced by the lambda
                co = fn.__code__.co_names
                name = next((n for n in co if n in ("step_prologue", "position_ids", "ln_fwd", "ln_bwd", "softmax_fwd",
                                                    "softmax_bwd", "attn_fwd", "attn_bwd", "head_fwd", "head_bwd",
                                                    "embed_bwd", "colsum", "zero_")), "op")
            plan.names.append(name)

        if train:
            o = self.opt
            sch = self.sched
            op(lambda: K.step_prologue(st, sch.base_lr, sch.kind, sch.total_steps, sch.warmup_steps, o.beta_1, o.beta_2,
                                       o.epsilon, o.weight_decay))
            self.dp.after_step_prologue(plan, B * S)
        op(lambda: K.position_ids(ids, pos, cfg.model_type == "roberta", cfg.pad_token_id))
        type0 = self.W("emb.tok_type.weight")[0] if cfg.type_vocab_size > 0 else None

        def X(i):  # input of layer i
            if train:
                return x_in[i]
            return xa if i % 2 == 0 else xb

        op(lambda: K.ln_fwd(y=X(0), pre=pre0, mean=mean0, rstd=rstd0, gamma=self.W("emb.ln.weight"),
                            beta=self.W("emb.ln.bias"), eps=eps, ids=ids.view(-1), pos_ids=pos.view(-1),
                            word=self.W("emb.word.weight"), pos=self.W("emb.pos.weight"), type0=type0, state=sref,
                            p_out=p_hid, site_out=9000))
        for i in range(L):
            j = i if train else 0
            pfx = f"layers.{i}."
            x, xn = X(i), X(i + 1)
            self._linear_fwd(plan, x, pfx + "qkv", qkv[j], M)
            if fused:
                op(lambda j=j, i=i: K.attn_fwd(qkv[j], ctx[j], lse[j], mask, B, H, S, scale, p_drop=p_att,
                                               site=i * 8 + 1, state=sref))
            else:
                # scores = scale * Q Kᵀ  per (head, batch)
                self._gemm_op(plan, A=qkv[j], B=(qkv[j], d), C=P[j], lda=3 * d, ldb=3 * d, ldc=S, a_bs0=dh,
                              a_bs1=S * 3 * d, b_bs0=dh, b_bs1=S * 3 * d, c_bs0=S * S, c_bs1=H * S * S, M=S, N=S, K=dh,
                              nb0=H, nb1=B, alpha=scale)
                op(lambda j=j, i=i: K.softmax_fwd(P[j], mask, pd=(Pd[j] if p_att > 0 else None), p_drop=p_att,
                                                  site=i * 8 + 1, state=sref))
                # ctx = Pd · V   (V read MN-major straight out of the fused QKV buffer)
                self._gemm_op(plan, A=Pd[j], B=(qkv[j], 2 * d), C=ctx[j], lda=S, ldb=3 * d, ldc=d, a_bs0=S * S,
                              a_bs1=H * S * S, b_bs0=dh, b_bs1=S * 3 * d, c_bs0=dh, c_bs1=S * d, M=S, N=dh, K=S,
                              nb0=H, nb1=B, b_t=True)
            self._linear_fwd(plan, ctx[j], pfx + "out", r1[j], M)
            op(lambda j=j, i=i, x=x: K.ln_fwd(a=r1[j], resid=x, pre=r1[j], y=h1[j], mean=mean1[j], rstd=rstd1[j],
                                              gamma=self.W(f"layers.{i}.ln1.weight"), beta=self.W(f"layers.{i}.ln1.bias"),
                                              eps=eps, state=sref, p_in=p_attn_out, site_in=i * 8 + 2))
            self._linear_fwd(plan, h1[j], pfx + "ffn1", f[j], M, epilogue=G.EPI_GELU, aux=u[j] if train else None)
            self._linear_fwd(plan, f[j], pfx + "ffn2", r2[j], M)
            op(lambda j=j, i=i, xn=xn: K.ln_fwd(a=r2[j], resid=h1[j], pre=r2[j], y=xn, mean=mean2[j], rstd=rstd2[j],
                                                gamma=self.W(f"layers.{i}.ln2.weight"),
                                                beta=self.W(f"layers.{i}.ln2.bias"), eps=eps, state=sref, p_in=p_hid,
                                                site_in=i * 8 + 3))
        x_last = X(L)
        C = cfg.num_labels
        logits = self._buf(plan, "logits", (B, C), f32)
        dlogits = self._buf(plan, "dlogits", (B, C), f32)
        hh, dz = self._buf(plan, "head_h", (B, d), f32), self._buf(plan, "head_dz", (B, d), f32)
        dx = self._buf(plan, "dx", (M, d)) if train else None
        hargs = K.make_head_args(
            x=x_last, w1=self.W("head.dense.weight"), b1=self.W("head.dense.bias"), w2=self.W("head.out.weight"),
            b2=self.W("head.out.bias"), labels=labels, logits=logits, dlogits=dlogits, h=hh, dz=dz, state=st, B=B, S=S,
            d=d, num_labels=C, kind=cfg.model_type, p_drop=p_cls, site0=9001, site1=9002,
            dw1=self.Gr("head.dense.weight") if train else None, db1=self.Gr("head.dense.bias") if train else None,
            dw2=self.Gr("head.out.weight") if train else None, db2=self.Gr("head.out.bias") if train else None, dx=dx)
        plan.head_args = hargs
        op(lambda: K.head_fwd(hargs, train), 2)
        if not train:
            return plan

        # ---------------------------------------------------------------- backward
        op(lambda: K.head_bwd(hargs), 4)
        bufA, bufB, bufC, bufD = (self._buf(plan, n, (M, d)) for n in ("gA", "gB", "gC", "gD"))
        dctx = self._buf(plan, "dctx", (M, d))
        du = self._buf(plan, "du", (M, F))
        dP = None if fused else self._buf(plan, "dP", (B, H, S, S))
        attn_ws = self._shared("attn_ws", K.attn_bwd_workspace_bytes(B, H, S), torch.uint8) if fused else None
        dqkv = self._buf(plan, "dqkv", (M, 3 * d))
        lnp = self._scratch.get("ln_acc")  # fp32 column accumulator of the LayerNorm backward: zero once, kept zeroed
        if lnp is None:
            lnp = self._scratch["ln_acc"] = torch.zeros(K.ln_bwd_partials_floats(d), device=self.device)
        self.dp.begin_backward(plan)
        for i in reversed(range(L)):
            pfx = f"layers.{i}."
            # LN2 backward: dx -> d(r2) in A ; dropout' -> B
            da2 = bufB if p_hid > 0 else None
            op(lambda i=i, da2=da2: K.ln_bwd(dy=dx, pre=r2[i], mean=mean2[i], rstd=rstd2[i],
                                             gamma=self.W(f"layers.{i}.ln2.weight"), dpre=bufA, da=da2,
                                             dgamma=self.Gr(f"layers.{i}.ln2.weight"), dbeta=self.Gr(f"layers.{i}.ln2.bias"),
                                             dbias=self.Gr(f"layers.{i}.ffn2.bias"), partials=lnp, state=st, p_in=p_hid,
                                             site_in=i * 8 + 3), 1)
            dg = da2 if da2 is not None else bufA
            # FFN2: dW2, du = (dg·W2) ∘ gelu'(u)
            self._linear_bwd(plan, dg, f[i], pfx + "ffn2", du, M, epilogue=G.EPI_GELU_BWD, aux=u[i], ldaux=F,
                             bias_grad=False)
            # FFN1: dW1, db1, dh1 = du·W1 + d(r2)
            self._linear_bwd(plan, du, h1[i], pfx + "ffn1", bufC, M, epilogue=G.EPI_ADD, aux=bufA, ldaux=d)
            if self.dp.local_early_buckets() or dg is bufA:
                # without hidden dropout (dg = bufA) the FFN2 weight-gradient GEMM on the side stream reads
                # bufA, which the LN1 backward below overwrites (dpre = bufA).  (A distributed bucket released
                # here does not need the join: its kernel waits for the side stream from the comm stream.)
                self._join_side(plan)
            self.dp.after_grads(plan, [pfx + n for n in ("ln2.weight", "ln2.bias", "ffn2.weight", "ffn2.bias",
                                                          "ffn1.weight", "ffn1.bias")])
            # LN1 backward: dh1 (C) -> d(r1) in A ; dropout' -> B
            da1 = bufD if p_attn_out > 0 else None  # (bufB may still be read by the FFN2 wgrad on the side stream)
            op(lambda i=i, da1=da1: K.ln_bwd(dy=bufC, pre=r1[i], mean=mean1[i], rstd=rstd1[i],
                                             gamma=self.W(f"layers.{i}.ln1.weight"), dpre=bufA, da=da1,
                                             dgamma=self.Gr(f"layers.{i}.ln1.weight"), dbeta=self.Gr(f"layers.{i}.ln1.bias"),
                                             dbias=self.Gr(f"layers.{i}.out.bias"), partials=lnp, state=st,
                                             p_in=p_attn_out, site_in=i * 8 + 2), 1)
            da = da1 if da1 is not None else bufA
            self._linear_bwd(plan, da, ctx[i], pfx + "out", dctx, M, bias_grad=False)
            if fused:
                op(lambda i=i: K.attn_bwd(qkv[i], ctx[i], dctx, lse[i], mask, dqkv, attn_ws, B, H, S, scale,
                                          p_drop=p_att, site=i * 8 + 1, state=st), 3)
            else:
                self._attention_bwd_unfused(plan, i, qkv, Pd, P, dP, dctx, dqkv, B, S, H, dh, d, scale, p_att, st, op)
            # QKV projection: dWqkv, dbqkv, dx = dqkv·Wqkv + d(r1)
            self._linear_bwd(plan, dqkv, x_in[i], pfx + "qkv", dx, M, epilogue=G.EPI_ADD, aux=bufA, ldaux=d)
            # the next layer overwrites the gradient buffers the side stream is still reading
            self._join_side(plan)
            self.dp.after_grads(plan, [pfx + n for n in ("ln1.weight", "ln1.bias", "out.weight", "out.bias",
                                                          "qkv.weight", "qkv.bias")])
        # embeddings
        op(lambda: K.ln_bwd(dy=dx, pre=pre0, mean=mean0, rstd=rstd0, gamma=self.W("emb.ln.weight"), dpre=bufA,
                            dgamma=self.Gr("emb.ln.weight"), dbeta=self.Gr("emb.ln.bias"), partials=lnp, state=st,
                            p_out=p_hid, site_out=9000), 1)
        wg, pg = self.Gr("emb.word.weight"), self.Gr("emb.pos.weight")
        if self.dp.sparse_embedding():
            # only the rows that are live in this step are ever read: the marking kernel cleared them long ago
            op(lambda: pg.zero_(), 1)
            self.dp.before_embedding_scatter(plan)
        else:
            op(lambda: (wg.zero_(), pg.zero_()), 2)
        op(lambda: K.embed_bwd(bufA, ids.view(-1), pos.view(-1), wg, pg))
        if cfg.type_vocab_size > 0:
            tg = self.Gr("emb.tok_type.weight")
            part = self._shared("colsum_partials", K.colsum_partials_floats(max(3 * d, F)))
            op(lambda: (tg.zero_(), K.colsum(bufA, tg[0], part)), 3)
        names = ["emb.ln.weight", "emb.ln.bias", "emb.pos.weight", "emb.word.weight"]
        if cfg.type_vocab_size > 0:
            names.append("emb.tok_type.weight")
        self._join_side(plan)
        self.dp.after_grads(plan, names)
        self.dp.end_backward(plan)
        return plan

    # ------------------------------------------------------------------ execution
    def _get_plan(self, B: int, S: int, train: bool) -> Plan:
        key = (B, S, train)
        plan = self.plans.get(key)
        if plan is None:
            # size the shared scratch before recording so pointers stay fixed
            self._shared("colsum_partials", K.colsum_partials_floats(max(3 * self.cfg.hidden_size,
                                                                         self.cfg.intermediate_size)))
            plan = self.build_plan(B, S, train)
            self.plans[key] = plan
        return plan

    def _run(self, plan: Plan) -> None:
        if self.use_cuda_graph and plan.graph is not None:
            plan.graph.replay()
            return
        if os.environ.get("HSTD_NVTX") == "1":
            # eager, one NVTX range per recorded op: readable timelines in Nsight
            from ..utils.timers import nvtx_range

            names = plan.names + ["dp/optimizer"] * (len(plan.ops) - len(plan.names))
            for f, n in zip(plan.ops, names):
                with nvtx_range(n):
                    f()
            return
        for f in plan.ops:
            f()
        plan.warm += 1
        if self.use_cuda_graph and plan.warm >= 2 and plan.graph is None and self.dp.capturable():
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for f in plan.ops:
                    f()
            plan.graph = g
            # capture executes nothing; the ops above already ran eagerly for this step

    def _stage_inputs(self, plan: Plan, feats, labels) -> None:
        plan.bufs["ids"].copy_(feats["input_ids"], non_blocking=True)
        plan.bufs["mask"].copy_(feats["attention_mask"], non_blocking=True)
        plan.bufs["labels"].copy_(labels, non_blocking=True)

    def train_step(self, feats, labels):
        assert self.opt is not None, "call compile()/fit() first"
        B, S = feats["input_ids"].shape
        plan = self._get_plan(B, S, True)
        self._stage_inputs(plan, feats, labels)
        self._run(plan)
        return None  # metrics accumulate on the device, see read_metrics()

    def eval_step(self, feats, labels):
        B, S = feats["input_ids"].shape
        plan = self._get_plan(B, S, False)
        self._stage_inputs(plan, feats, labels)
        self._run(plan)
        return None

    def last_step_metrics(self) -> torch.Tensor:
        """[loss_sum, correct, count] of the most recent step (device tensor, fp32)."""
        return self.state[METRICS_OFFSET:METRICS_OFFSET + 12].view(torch.float32)

    def read_metrics(self) -> Dict[str, float]:
        l, c, n, _ = self._metrics_view.tolist()  # one D2H read
        n = max(n, 1.0)
        return {"loss": l / n, "sparse_categorical_accuracy": c / n, "count": n}

    def reset_metrics(self) -> None:
        K.metrics_reset(self.state)

    def logits(self, B: int, S: int, train: bool = False) -> torch.Tensor:
        return self.plans[(B, S, train)].bufs["logits"]

    def launches_per_step(self, B: int, S: int, train: bool = True) -> int:
        return self._get_plan(B, S, train).launches

    def profile(self, B: int, S: int, train: bool = True, repeats: int = 5):
        """Per-op device time of one step (eager, CUDA events): [(name, calls, us)]."""
        from ..utils.timers import profile_ops

        plan = self._get_plan(B, S, train)
        names = plan.names + ["dp/optimizer"] * (len(plan.ops) - len(plan.names))
        return profile_ops(plan.ops, names, repeats)

    def poll_health(self) -> None:
        """Cheap (non-blocking) check of the all-reduce watchdog; raises if a peer went missing."""
        self.dp.poll_errors()

    def finish(self) -> None:
        torch.cuda.synchronize()
        self.dp.check_errors()
