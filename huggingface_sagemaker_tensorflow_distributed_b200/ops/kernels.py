"""ctypes front-ends of the non-GEMM kernels (csrc/norm_act.cu, head_loss.cu, optimizer.cu).

Thin by design: the engine pre-allocates every buffer and calls these with raw tensors; the
wrappers only marshal pointers.  The GPU tests (tests/test_kernels_gpu.py, test_attention_gpu.py)
compare each against a plain PyTorch fp32 statement of the same math written next to the test.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib

SCHED = {"constant": 0, "linear": 1}
HEAD_KIND = {"distilbert": 0, "bert": 1, "roberta": 2}


class StepState(C.Structure):
    _fields_ = [("step", C.c_int64), ("seed", C.c_uint64), ("lr", C.c_float), ("lr_t", C.c_float),
                ("decay", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
                ("metrics", C.c_float * 4), ("metrics_acc", C.c_double * 4)]


STEP_STATE_BYTES = C.sizeof(StepState)


class LnArgs(C.Structure):
    _fields_ = [("a", C.c_void_p), ("resid", C.c_void_p), ("gamma", C.c_void_p), ("beta", C.c_void_p),
                ("pre", C.c_void_p), ("y", C.c_void_p), ("mean", C.c_void_p), ("rstd", C.c_void_p),
                ("ids", C.c_void_p), ("pos_ids", C.c_void_p), ("word", C.c_void_p), ("pos", C.c_void_p),
                ("type0", C.c_void_p), ("state", C.c_void_p), ("M", C.c_int32), ("d", C.c_int32),
                ("eps", C.c_float), ("p_in", C.c_float), ("p_out", C.c_float), ("site_in", C.c_uint32),
                ("site_out", C.c_uint32)]


class LnBwdArgs(C.Structure):
    _fields_ = [("dy", C.c_void_p), ("pre", C.c_void_p), ("mean", C.c_void_p), ("rstd", C.c_void_p),
                ("gamma", C.c_void_p), ("dpre", C.c_void_p), ("da", C.c_void_p), ("dgamma", C.c_void_p),
                ("dbeta", C.c_void_p), ("dbias", C.c_void_p), ("partials", C.c_void_p), ("state", C.c_void_p),
                ("M", C.c_int32), ("d", C.c_int32), ("p_in", C.c_float), ("p_out", C.c_float),
                ("site_in", C.c_uint32), ("site_out", C.c_uint32)]


class HeadArgs(C.Structure):
    _fields_ = [("x", C.c_void_p), ("w1", C.c_void_p), ("b1", C.c_void_p), ("w2", C.c_void_p), ("b2", C.c_void_p),
                ("labels", C.c_void_p), ("logits", C.c_void_p), ("dlogits", C.c_void_p), ("h", C.c_void_p),
                ("dz", C.c_void_p), ("dw1", C.c_void_p), ("db1", C.c_void_p), ("dw2", C.c_void_p),
                ("db2", C.c_void_p), ("dx", C.c_void_p), ("state", C.c_void_p), ("B", C.c_int32), ("S", C.c_int32),
                ("d", C.c_int32), ("C", C.c_int32), ("kind", C.c_int32), ("train", C.c_int32),
                ("p_drop", C.c_float), ("site0", C.c_uint32), ("site1", C.c_uint32)]


@_lib.registrar
def _register(lib) -> None:
    vp, i32, i64, f32, u32, u64 = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_uint32, C.c_uint64
    lib.hstd_step_state_init.argtypes = [vp, u64, i64, vp]
    lib.hstd_step_prologue.argtypes = [vp, f32, i32, i64, i64, f32, f32, f32, f32, vp]
    lib.hstd_metrics_reset.argtypes = [vp, vp]
    lib.hstd_adamw_flat.argtypes = [vp, vp, vp, vp, vp, vp, i64, f32, vp]
    lib.hstd_ln_fwd.argtypes = [C.POINTER(LnArgs), vp]
    lib.hstd_ln_bwd.argtypes = [C.POINTER(LnBwdArgs), vp]
    lib.hstd_ln_bwd_partials_floats.argtypes = [i32]
    lib.hstd_ln_bwd_partials_floats.restype = i64
    lib.hstd_colsum.argtypes = [vp, i64, i32, i32, vp, vp, vp]
    lib.hstd_colsum_partials_floats.argtypes = [i32]
    lib.hstd_colsum_partials_floats.restype = i64
    lib.hstd_softmax_fwd.argtypes = [vp, vp, vp, i32, i32, i32, i32, f32, u32, vp, vp]
    lib.hstd_softmax_bwd.argtypes = [vp, vp, i32, i32, i32, i32, f32, u32, vp, vp]
    lib.hstd_attn_fwd.argtypes = [vp, vp, vp, vp, i32, i32, i32, f32, f32, u32, vp, vp]
    lib.hstd_attn_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, f32, f32, u32, vp, vp]
    lib.hstd_attn_bwd_workspace_bytes.argtypes = [i32, i32, i32]
    lib.hstd_attn_bwd_workspace_bytes.restype = i64
    lib.hstd_position_ids.argtypes = [vp, vp, i32, i32, i32, i32, vp]
    lib.hstd_embed_bwd.argtypes = [vp, vp, vp, vp, vp, i32, i32, vp]
    lib.hstd_head_fwd.argtypes = [C.POINTER(HeadArgs), vp]
    lib.hstd_head_bwd.argtypes = [C.POINTER(HeadArgs), vp]
    for n in ("hstd_step_state_init", "hstd_step_prologue", "hstd_metrics_reset", "hstd_adamw_flat", "hstd_ln_fwd",
              "hstd_ln_bwd", "hstd_colsum", "hstd_softmax_fwd", "hstd_softmax_bwd", "hstd_position_ids",
              "hstd_embed_bwd", "hstd_head_fwd", "hstd_head_bwd", "hstd_attn_fwd", "hstd_attn_bwd"):
        getattr(lib, n).restype = C.c_int


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _s(stream=None):
    return _lib.stream_ptr(stream)


# ----------------------------------------------------------------------------- step state
def new_step_state(device, seed: int = 0, step: int = 0) -> torch.Tensor:
    st = torch.zeros(STEP_STATE_BYTES, dtype=torch.uint8, device=device)
    _lib.check(_lib.load().hstd_step_state_init(st.data_ptr(), seed, step, _s()), "step_state_init")
    return st


def read_step_state(st: torch.Tensor) -> StepState:
    raw = bytes(st.cpu().numpy().tobytes())
    return StepState.from_buffer_copy(raw)


def step_prologue(st: torch.Tensor, base_lr: float, schedule: str, total_steps: int, warmup_steps: int, beta1: float,
                  beta2: float, eps: float, weight_decay: float, stream=None) -> None:
    _lib.check(_lib.load().hstd_step_prologue(st.data_ptr(), base_lr, SCHED[schedule], total_steps, warmup_steps,
                                              beta1, beta2, eps, weight_decay, _s(stream)), "step_prologue")


def metrics_reset(st: torch.Tensor, stream=None) -> None:
    _lib.check(_lib.load().hstd_metrics_reset(st.data_ptr(), _s(stream)), "metrics_reset")


def adamw_flat(st, grads, master, m, v, w_bf16, grad_scale: float = 1.0, stream=None) -> None:
    _lib.check(_lib.load().hstd_adamw_flat(st.data_ptr(), grads.data_ptr(), master.data_ptr(), m.data_ptr(),
                                           v.data_ptr(), w_bf16.data_ptr(), grads.numel(), grad_scale, _s(stream)),
               "adamw_flat")


# ----------------------------------------------------------------------------- LayerNorm
def ln_fwd(*, y, gamma, beta, eps, a=None, resid=None, pre=None, mean=None, rstd=None, ids=None, pos_ids=None,
           word=None, pos=None, type0=None, state=None, p_in=0.0, p_out=0.0, site_in=0, site_out=0, stream=None):
    g = LnArgs()
    g.a, g.resid, g.gamma, g.beta, g.pre, g.y = _p(a), _p(resid), _p(gamma), _p(beta), _p(pre), _p(y)
    g.mean, g.rstd, g.ids, g.pos_ids = _p(mean), _p(rstd), _p(ids), _p(pos_ids)
    g.word, g.pos, g.type0, g.state = _p(word), _p(pos), _p(type0), _p(state)
    g.M, g.d = y.shape[0], y.shape[1]
    g.eps, g.p_in, g.p_out, g.site_in, g.site_out = eps, p_in, p_out, site_in, site_out
    _lib.check(_lib.load().hstd_ln_fwd(C.byref(g), _s(stream)), "ln_fwd")


def ln_bwd_partials_floats(d: int) -> int:
    return _lib.load().hstd_ln_bwd_partials_floats(d)


def ln_bwd(*, dy, pre, mean, rstd, gamma, dpre, dgamma, dbeta, partials, da=None, dbias=None, state=None, p_in=0.0,
           p_out=0.0, site_in=0, site_out=0, stream=None):
    g = LnBwdArgs()
    g.dy, g.pre, g.mean, g.rstd, g.gamma = _p(dy), _p(pre), _p(mean), _p(rstd), _p(gamma)
    g.dpre, g.da, g.dgamma, g.dbeta, g.dbias = _p(dpre), _p(da), _p(dgamma), _p(dbeta), _p(dbias)
    g.partials, g.state = _p(partials), _p(state)
    g.M, g.d = dy.shape[0], dy.shape[1]
    g.p_in, g.p_out, g.site_in, g.site_out = p_in, p_out, site_in, site_out
    _lib.check(_lib.load().hstd_ln_bwd(C.byref(g), _s(stream)), "ln_bwd")


def colsum_partials_floats(n: int) -> int:
    return _lib.load().hstd_colsum_partials_floats(n)


def colsum(x, out, partials, stream=None):
    _lib.check(_lib.load().hstd_colsum(x.data_ptr(), x.stride(0), x.shape[0], x.shape[1], out.data_ptr(),
                                       partials.data_ptr(), _s(stream)), "colsum")


# ----------------------------------------------------------------------------- softmax
def softmax_fwd(scores, key_mask, pd=None, p_drop=0.0, site=0, state=None, stream=None):
    B, H, Sq, Sk = scores.shape
    _lib.check(_lib.load().hstd_softmax_fwd(scores.data_ptr(), _p(pd), key_mask.data_ptr(), B, H, Sq, Sk, p_drop, site,
                                            _p(state), _s(stream)), "softmax_fwd")


def softmax_bwd(p, dpd, p_drop=0.0, site=0, state=None, stream=None):
    B, H, Sq, Sk = p.shape
    _lib.check(_lib.load().hstd_softmax_bwd(p.data_ptr(), dpd.data_ptr(), B, H, Sq, Sk, p_drop, site, _p(state),
                                            _s(stream)), "softmax_bwd")


# ----------------------------------------------------------------------------- fused attention
def attn_fwd(qkv, ctx, lse, key_mask, B, H, S, scale, p_drop=0.0, site=0, state=None, stream=None):
    _lib.check(_lib.load().hstd_attn_fwd(qkv.data_ptr(), ctx.data_ptr(), _p(lse), key_mask.data_ptr(), B, H, S, scale,
                                         p_drop, site, _p(state), _s(stream)), "attn_fwd")


def attn_bwd_workspace_bytes(B, H, S) -> int:
    return _lib.load().hstd_attn_bwd_workspace_bytes(B, H, S)


def attn_bwd(qkv, ctx, dctx, lse, key_mask, dqkv, workspace, B, H, S, scale, p_drop=0.0, site=0, state=None,
             stream=None):
    _lib.check(_lib.load().hstd_attn_bwd(qkv.data_ptr(), ctx.data_ptr(), dctx.data_ptr(), lse.data_ptr(),
                                         key_mask.data_ptr(), dqkv.data_ptr(), workspace.data_ptr(), B, H, S, scale,
                                         p_drop, site, _p(state), _s(stream)), "attn_bwd")


# ----------------------------------------------------------------------------- embeddings
def position_ids(ids, out, roberta: bool, pad_id: int, stream=None):
    B, S = ids.shape
    _lib.check(_lib.load().hstd_position_ids(ids.data_ptr(), out.data_ptr(), B, S, 1 if roberta else 0, pad_id,
                                             _s(stream)), "position_ids")


def embed_bwd(de, ids, pos_ids, word_grad, pos_grad, stream=None):
    _lib.check(_lib.load().hstd_embed_bwd(de.data_ptr(), ids.data_ptr(), pos_ids.data_ptr(), word_grad.data_ptr(),
                                          pos_grad.data_ptr(), de.shape[0], de.shape[1], _s(stream)), "embed_bwd")


# ----------------------------------------------------------------------------- head
def make_head_args(*, x, w1, b1, w2, b2, labels, logits, dlogits, h, dz, state, B, S, d, num_labels, kind, p_drop,
                   site0, site1, dw1=None, db1=None, dw2=None, db2=None, dx=None) -> HeadArgs:
    g = HeadArgs()
    g.x, g.w1, g.b1, g.w2, g.b2, g.labels = _p(x), _p(w1), _p(b1), _p(w2), _p(b2), _p(labels)
    g.logits, g.dlogits, g.h, g.dz = _p(logits), _p(dlogits), _p(h), _p(dz)
    g.dw1, g.db1, g.dw2, g.db2, g.dx, g.state = _p(dw1), _p(db1), _p(dw2), _p(db2), _p(dx), _p(state)
    g.B, g.S, g.d, g.C, g.kind, g.train = B, S, d, num_labels, HEAD_KIND[kind] if isinstance(kind, str) else kind, 1
    g.p_drop, g.site0, g.site1 = p_drop, site0, site1
    return g


def head_fwd(args: HeadArgs, train: bool, stream=None):
    args.train = 1 if train else 0
    _lib.check(_lib.load().hstd_head_fwd(C.byref(args), _s(stream)), "head_fwd")


def head_bwd(args: HeadArgs, stream=None):
    args.train = 1
    _lib.check(_lib.load().hstd_head_bwd(C.byref(args), _s(stream)), "head_bwd")
