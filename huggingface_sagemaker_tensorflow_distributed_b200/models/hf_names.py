"""Canonical parameter names ↔ Hugging Face checkpoint names.

Canonical names (fused QKV) are what the reference module and the B200 engine use; the HF
names are what ``save_pretrained`` must emit so that ``transformers`` —and through
``from_pt=True`` the TF classes the reference uses (scripts/train.py:117)— can reload the
checkpoint.  The mapping is one-to-one except ``layers.i.qkv`` ↔ three q/k/v tensors.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import torch

from .configuration import EncoderConfig


def _layer_table(cfg: EncoderConfig, i: int) -> List[Tuple[str, str]]:
    if cfg.model_type == "distilbert":
        p = f"distilbert.transformer.layer.{i}."
        return [("out", p + "attention.out_lin"), ("ln1", p + "sa_layer_norm"), ("ffn1", p + "ffn.lin1"),
                ("ffn2", p + "ffn.lin2"), ("ln2", p + "output_layer_norm")]
    p = f"{cfg.model_type}.encoder.layer.{i}."
    return [("out", p + "attention.output.dense"), ("ln1", p + "attention.output.LayerNorm"),
            ("ffn1", p + "intermediate.dense"), ("ffn2", p + "output.dense"), ("ln2", p + "output.LayerNorm")]


def _qkv_names(cfg: EncoderConfig, i: int) -> List[str]:
    if cfg.model_type == "distilbert":
        p = f"distilbert.transformer.layer.{i}.attention."
        return [p + "q_lin", p + "k_lin", p + "v_lin"]
    p = f"{cfg.model_type}.encoder.layer.{i}.attention.self."
    return [p + "query", p + "key", p + "value"]


def _static_table(cfg: EncoderConfig) -> List[Tuple[str, str]]:
    mt = cfg.model_type
    e = f"{mt}.embeddings."
    t = [("emb.word.weight", e + "word_embeddings.weight"), ("emb.pos.weight", e + "position_embeddings.weight"),
         ("emb.ln.weight", e + "LayerNorm.weight"), ("emb.ln.bias", e + "LayerNorm.bias")]
    if cfg.type_vocab_size > 0:
        t.append(("emb.tok_type.weight", e + "token_type_embeddings.weight"))
    if mt == "distilbert":
        t += [("head.dense.weight", "pre_classifier.weight"), ("head.dense.bias", "pre_classifier.bias"),
              ("head.out.weight", "classifier.weight"), ("head.out.bias", "classifier.bias")]
    elif mt == "bert":
        t += [("head.dense.weight", "bert.pooler.dense.weight"), ("head.dense.bias", "bert.pooler.dense.bias"),
              ("head.out.weight", "classifier.weight"), ("head.out.bias", "classifier.bias")]
    else:
        t += [("head.dense.weight", "classifier.dense.weight"), ("head.dense.bias", "classifier.dense.bias"),
              ("head.out.weight", "classifier.out_proj.weight"), ("head.out.bias", "classifier.out_proj.bias")]
    return t


def canonical_to_hf(cfg: EncoderConfig, sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    out: Dict[str, torch.Tensor] = {}
    for c, h in _static_table(cfg):
        out[h] = sd[c]
    d = cfg.hidden_size
    for i in range(cfg.num_layers):
        for c, h in _layer_table(cfg, i):
            for suf in ("weight", "bias"):
                out[f"{h}.{suf}"] = sd[f"layers.{i}.{c}.{suf}"]
        w, b = sd[f"layers.{i}.qkv.weight"], sd[f"layers.{i}.qkv.bias"]
        for j, h in enumerate(_qkv_names(cfg, i)):
            out[h + ".weight"] = w[j * d:(j + 1) * d].clone()
            out[h + ".bias"] = b[j * d:(j + 1) * d].clone()
    return out


def hf_to_canonical(cfg: EncoderConfig, hf: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    out: Dict[str, torch.Tensor] = {}
    missing: List[str] = []

    def take(name: str):
        if name not in hf:
            missing.append(name)
            return None
        return hf[name]

    for c, h in _static_table(cfg):
        v = take(h)
        if v is not None:
            out[c] = v
    for i in range(cfg.num_layers):
        for c, h in _layer_table(cfg, i):
            for suf in ("weight", "bias"):
                v = take(f"{h}.{suf}")
                if v is not None:
                    out[f"layers.{i}.{c}.{suf}"] = v
        ws = [take(h + ".weight") for h in _qkv_names(cfg, i)]
        bs = [take(h + ".bias") for h in _qkv_names(cfg, i)]
        if all(w is not None for w in ws) and all(b is not None for b in bs):
            out[f"layers.{i}.qkv.weight"] = torch.cat(ws, 0)
            out[f"layers.{i}.qkv.bias"] = torch.cat(bs, 0)
    # A checkpoint of the bare encoder (e.g. a masked-LM pretrain) has no classifier: that is
    # the normal fine-tuning start, the head stays randomly initialised.
    head_missing = [m for m in missing if "classifier" in m or "pooler" in m]
    hard_missing = [m for m in missing if m not in head_missing]
    if hard_missing:
        raise KeyError(f"checkpoint is missing {len(hard_missing)} tensors, e.g. {hard_missing[:4]}")
    return out


def canonical_names(cfg: EncoderConfig) -> List[str]:
    names = [c for c, _ in _static_table(cfg)]
    for i in range(cfg.num_layers):
        for c in ("qkv", "out", "ln1", "ffn1", "ffn2", "ln2"):
            names += [f"layers.{i}.{c}.weight", f"layers.{i}.{c}.bias"]
    return names
