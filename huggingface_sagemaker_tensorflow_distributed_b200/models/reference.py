"""Plain-PyTorch encoder classifier: the numerical oracle and the CPU execution path.

One module covers the three families; the math follows the HF definitions the reference
loads:

  * DistilBERT: word+pos → LN → drop; L×[MHA → LN(x+·) → FFN(gelu) → drop → LN(x+·)];
    head ``h[:,0] → Linear → ReLU → drop(0.2) → Linear``
  * BERT: + token-type embedding (ids are all 0 because the reference drops
    ``token_type_ids``, scripts/train.py:78-83), dropout after the attention output
    projection, head ``tanh(Linear(h[:,0])) → drop → Linear``
  * RoBERTa: position ids ``cumsum(ids != pad) + pad``, LN eps 1e-5, head
    ``drop → Linear → tanh → drop → Linear``

Parameters use a canonical, engine-friendly naming (fused QKV); ``hf_names.py`` maps them
one-to-one onto the Hugging Face checkpoint names so ``save_pretrained`` output reloads
into ``transformers`` (and hence ``TFAutoModel...from_pretrained(dir, from_pt=True)``).
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .configuration import EncoderConfig


class _Layer(nn.Module):
    def __init__(self, cfg: EncoderConfig):
        super().__init__()
        d, f = cfg.hidden_size, cfg.intermediate_size
        self.qkv = nn.Linear(d, 3 * d)
        self.out = nn.Linear(d, d)
        self.ln1 = nn.LayerNorm(d, eps=cfg.layer_norm_eps)
        self.ffn1 = nn.Linear(d, f)
        self.ffn2 = nn.Linear(f, d)
        self.ln2 = nn.LayerNorm(d, eps=cfg.layer_norm_eps)


class _Embeddings(nn.Module):
    def __init__(self, cfg: EncoderConfig):
        super().__init__()
        self.word = nn.Embedding(cfg.vocab_size, cfg.hidden_size)
        self.pos = nn.Embedding(cfg.max_position_embeddings, cfg.hidden_size)
        if cfg.type_vocab_size > 0:
            self.tok_type = nn.Embedding(cfg.type_vocab_size, cfg.hidden_size)
        self.ln = nn.LayerNorm(cfg.hidden_size, eps=cfg.layer_norm_eps)


class _Head(nn.Module):
    def __init__(self, cfg: EncoderConfig):
        super().__init__()
        self.dense = nn.Linear(cfg.hidden_size, cfg.hidden_size)
        self.out = nn.Linear(cfg.hidden_size, cfg.num_labels)


def position_ids_for(cfg: EncoderConfig, input_ids: torch.Tensor) -> torch.Tensor:
    """[B,S] position ids: arange for BERT/DistilBERT, pad-aware cumsum for RoBERTa."""
    B, S = input_ids.shape
    if cfg.model_type == "roberta":
        nonpad = (input_ids != cfg.pad_token_id).to(torch.int64)
        return torch.cumsum(nonpad, dim=1) * nonpad + cfg.pad_token_id
    return torch.arange(S, device=input_ids.device, dtype=torch.int64).unsqueeze(0).expand(B, S)


class EncoderForSequenceClassification(nn.Module):
    def __init__(self, cfg: EncoderConfig, sdpa: bool = False):
        super().__init__()
        self.cfg = cfg
        self.sdpa = sdpa  # library fused attention (used by the DDP baseline only)
        self.emb = _Embeddings(cfg)
        self.layers = nn.ModuleList(_Layer(cfg) for _ in range(cfg.num_layers))
        self.head = _Head(cfg)
        self.reset_parameters()

    def reset_parameters(self) -> None:
        std = self.cfg.initializer_range
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.normal_(m.weight, mean=0.0, std=std)
                nn.init.zeros_(m.bias)
            elif isinstance(m, nn.Embedding):
                nn.init.normal_(m.weight, mean=0.0, std=std)
            elif isinstance(m, nn.LayerNorm):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)
        if self.cfg.model_type != "distilbert":
            with torch.no_grad():  # HF zeroes the padding row
                self.emb.word.weight[self.cfg.pad_token_id].zero_()

    def forward(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        cfg = self.cfg
        B, S = input_ids.shape
        H, dh = cfg.num_heads, cfg.head_dim
        if attention_mask is None:
            attention_mask = torch.ones_like(input_ids)
        x = self.emb.word(input_ids) + self.emb.pos(position_ids_for(cfg, input_ids))
        if cfg.type_vocab_size > 0:
            x = x + self.emb.tok_type.weight[0]
        x = F.dropout(self.emb.ln(x), cfg.hidden_dropout, self.training)
        neg = torch.finfo(x.dtype).min
        bias = torch.zeros(B, 1, 1, S, dtype=x.dtype, device=x.device).masked_fill(
            attention_mask[:, None, None, :] == 0, neg)
        post_attn_dropout = cfg.hidden_dropout if cfg.model_type != "distilbert" else 0.0
        for lyr in self.layers:
            qkv = lyr.qkv(x).view(B, S, 3, H, dh)
            q, k, v = (qkv[:, :, i].transpose(1, 2) for i in range(3))  # [B,H,S,dh]
            if self.sdpa:
                ctx = F.scaled_dot_product_attention(
                    q, k, v, attn_mask=bias, dropout_p=cfg.attention_dropout if self.training else 0.0)
                ctx = ctx.transpose(1, 2).reshape(B, S, H * dh)
            else:
                scores = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(dh) + bias
                probs = F.dropout(torch.softmax(scores, dim=-1), cfg.attention_dropout, self.training)
                ctx = torch.matmul(probs, v).transpose(1, 2).reshape(B, S, H * dh)
            a = F.dropout(lyr.out(ctx), post_attn_dropout, self.training)
            x = lyr.ln1(a + x)
            f = lyr.ffn2(F.gelu(lyr.ffn1(x)))
            x = lyr.ln2(F.dropout(f, cfg.hidden_dropout, self.training) + x)
        pooled = x[:, 0]
        if cfg.model_type == "distilbert":
            h = F.dropout(F.relu(self.head.dense(pooled)), cfg.classifier_dropout, self.training)
        elif cfg.model_type == "bert":
            h = F.dropout(torch.tanh(self.head.dense(pooled)), cfg.classifier_dropout, self.training)
        else:
            h = F.dropout(pooled, cfg.classifier_dropout, self.training)
            h = F.dropout(torch.tanh(self.head.dense(h)), cfg.classifier_dropout, self.training)
        return self.head.out(h)

    def num_parameters(self) -> int:
        return sum(p.numel() for p in self.parameters())
