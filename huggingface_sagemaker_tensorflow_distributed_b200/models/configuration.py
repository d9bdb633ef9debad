"""Encoder configurations for the three model families the reference fine-tunes through
``TFAutoModelForSequenceClassification.from_pretrained(name)`` (scripts/train.py:117,
scripts/singe_node_train.py:43): DistilBERT, BERT and RoBERTa.

There is no network here, so well-known hub names resolve to their published architecture
hyper-parameters (``KNOWN_MODELS``) with random-initialised weights; a local directory
written by ``save_pretrained`` resolves through its ``config.json``.
"""
from __future__ import annotations

import dataclasses
import json
import os
from dataclasses import dataclass
from typing import Any, Dict, Optional


@dataclass
class EncoderConfig:
    model_type: str = "distilbert"  # distilbert | bert | roberta
    vocab_size: int = 30522
    hidden_size: int = 768
    num_layers: int = 6
    num_heads: int = 12
    intermediate_size: int = 3072
    max_position_embeddings: int = 512
    type_vocab_size: int = 0  # 0 = no token-type table (DistilBERT)
    layer_norm_eps: float = 1e-12
    pad_token_id: int = 0
    hidden_dropout: float = 0.1  # embeddings / after attention-out / after FFN
    attention_dropout: float = 0.1
    classifier_dropout: float = 0.2  # DistilBERT seq_classif_dropout; BERT/RoBERTa: hidden dropout
    num_labels: int = 2
    initializer_range: float = 0.02
    model_max_length: int = 512  # tokenizer.model_max_length (scripts/train.py:81)
    name_or_path: str = ""

    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_heads

    # ---- HF config.json round trip -------------------------------------------------
    def to_hf_dict(self) -> Dict[str, Any]:
        common = {
            "vocab_size": self.vocab_size,
            "max_position_embeddings": self.max_position_embeddings,
            "initializer_range": self.initializer_range,
            "pad_token_id": self.pad_token_id,
            "id2label": {str(i): f"LABEL_{i}" for i in range(self.num_labels)},
            "label2id": {f"LABEL_{i}": i for i in range(self.num_labels)},
            "problem_type": "single_label_classification",
            "dtype": "float32",
            "torch_dtype": "float32",
            "transformers_version": "4.5.0+hstd_b200",
        }
        if self.model_type == "distilbert":
            return {
                **common,
                "architectures": ["DistilBertForSequenceClassification"],
                "model_type": "distilbert",
                "activation": "gelu",
                "dim": self.hidden_size,
                "n_layers": self.num_layers,
                "n_heads": self.num_heads,
                "hidden_dim": self.intermediate_size,
                "dropout": self.hidden_dropout,
                "attention_dropout": self.attention_dropout,
                "seq_classif_dropout": self.classifier_dropout,
                "qa_dropout": 0.1,
                "sinusoidal_pos_embds": False,
                "tie_weights_": True,
            }
        arch = "BertForSequenceClassification" if self.model_type == "bert" else "RobertaForSequenceClassification"
        d = {
            **common,
            "architectures": [arch],
            "model_type": self.model_type,
            "hidden_act": "gelu",
            "hidden_size": self.hidden_size,
            "num_hidden_layers": self.num_layers,
            "num_attention_heads": self.num_heads,
            "intermediate_size": self.intermediate_size,
            "hidden_dropout_prob": self.hidden_dropout,
            "attention_probs_dropout_prob": self.attention_dropout,
            "classifier_dropout": None,
            "type_vocab_size": self.type_vocab_size,
            "layer_norm_eps": self.layer_norm_eps,
            "position_embedding_type": "absolute",
        }
        if self.model_type == "roberta":
            d.update({"bos_token_id": 0, "eos_token_id": 2})
        return d

    @classmethod
    def from_hf_dict(cls, d: Dict[str, Any], name_or_path: str = "") -> "EncoderConfig":
        mt = d.get("model_type", "distilbert")
        num_labels = len(d["id2label"]) if "id2label" in d else int(d.get("num_labels", 2))
        if mt == "distilbert":
            return cls(model_type="distilbert", vocab_size=d.get("vocab_size", 30522), hidden_size=d.get("dim", 768),
                       num_layers=d.get("n_layers", 6), num_heads=d.get("n_heads", 12),
                       intermediate_size=d.get("hidden_dim", 3072),
                       max_position_embeddings=d.get("max_position_embeddings", 512), type_vocab_size=0,
                       layer_norm_eps=1e-12, pad_token_id=d.get("pad_token_id", 0),
                       hidden_dropout=d.get("dropout", 0.1), attention_dropout=d.get("attention_dropout", 0.1),
                       classifier_dropout=d.get("seq_classif_dropout", 0.2), num_labels=num_labels,
                       initializer_range=d.get("initializer_range", 0.02),
                       model_max_length=min(512, d.get("max_position_embeddings", 512)), name_or_path=name_or_path)
        if mt not in ("bert", "roberta"):
            raise ValueError(f"unsupported model_type {mt!r} (supported: distilbert, bert, roberta)")
        hd = d.get("hidden_dropout_prob", 0.1)
        cd = d.get("classifier_dropout")
        mpe = d.get("max_position_embeddings", 512)
        pad = d.get("pad_token_id", 1 if mt == "roberta" else 0)
        return cls(model_type=mt, vocab_size=d.get("vocab_size", 30522), hidden_size=d.get("hidden_size", 768),
                   num_layers=d.get("num_hidden_layers", 12), num_heads=d.get("num_attention_heads", 12),
                   intermediate_size=d.get("intermediate_size", 3072), max_position_embeddings=mpe,
                   type_vocab_size=d.get("type_vocab_size", 2 if mt == "bert" else 1),
                   layer_norm_eps=d.get("layer_norm_eps", 1e-12 if mt == "bert" else 1e-5), pad_token_id=pad,
                   hidden_dropout=hd, attention_dropout=d.get("attention_probs_dropout_prob", 0.1),
                   classifier_dropout=hd if cd is None else cd, num_labels=num_labels,
                   initializer_range=d.get("initializer_range", 0.02),
                   model_max_length=(mpe - pad - 1) if mt == "roberta" else min(512, mpe), name_or_path=name_or_path)

    def replace(self, **kw) -> "EncoderConfig":
        return dataclasses.replace(self, **kw)


def _distil(**kw):
    return EncoderConfig(model_type="distilbert", **kw)


def _bert(h, L, H, F, **kw):
    return EncoderConfig(model_type="bert", hidden_size=h, num_layers=L, num_heads=H, intermediate_size=F,
                         type_vocab_size=2, layer_norm_eps=1e-12, classifier_dropout=0.1, **kw)


def _roberta(h, L, H, F, **kw):
    return EncoderConfig(model_type="roberta", vocab_size=50265, hidden_size=h, num_layers=L, num_heads=H,
                         intermediate_size=F, max_position_embeddings=514, type_vocab_size=1, layer_norm_eps=1e-5,
                         pad_token_id=1, classifier_dropout=0.1, model_max_length=512, **kw)


KNOWN_MODELS = {
    "distilbert-base-uncased": lambda: _distil(),
    "distilbert-base-cased": lambda: _distil(vocab_size=28996),
    "bert-base-uncased": lambda: _bert(768, 12, 12, 3072),
    "bert-base-cased": lambda: _bert(768, 12, 12, 3072, vocab_size=28996),
    "bert-large-uncased": lambda: _bert(1024, 24, 16, 4096),
    "bert-large-cased": lambda: _bert(1024, 24, 16, 4096, vocab_size=28996),
    # the checked-in workload of the reference (launch.py:17)
    "bert-large-uncased-whole-word-masking": lambda: _bert(1024, 24, 16, 4096),
    "roberta-base": lambda: _roberta(768, 12, 12, 3072),
    "roberta-large": lambda: _roberta(1024, 24, 16, 4096),
    "distilroberta-base": lambda: _roberta(768, 6, 12, 3072),
    # tiny variants for tests / smoke runs
    "tiny-distilbert": lambda: _distil(vocab_size=1024, hidden_size=128, num_layers=2, num_heads=2,
                                       intermediate_size=256),
    "tiny-bert": lambda: _bert(128, 2, 2, 256, vocab_size=1024),
    "tiny-roberta": lambda: _roberta(128, 2, 2, 256).replace(vocab_size=1024),
    # smallest shapes the sm_100a kernels accept (hidden % 256 == 0, head_dim == 64)
    "tiny-distilbert-gpu": lambda: _distil(vocab_size=2048, hidden_size=256, num_layers=2, num_heads=4,
                                           intermediate_size=512),
    "tiny-bert-gpu": lambda: _bert(256, 2, 4, 512, vocab_size=2048),
    "tiny-roberta-gpu": lambda: _roberta(256, 2, 4, 512).replace(vocab_size=2048),
}


def resolve_config(name_or_path: str, num_labels: Optional[int] = None) -> EncoderConfig:
    """Hub name or local ``save_pretrained`` directory → EncoderConfig."""
    if name_or_path is None:
        raise ValueError("--model_name_or_path is required (the reference gives it no default, scripts/train.py:42)")
    cfg_file = os.path.join(name_or_path, "config.json")
    if os.path.isdir(name_or_path) and os.path.exists(cfg_file):
        with open(cfg_file) as f:
            cfg = EncoderConfig.from_hf_dict(json.load(f), name_or_path=name_or_path)
    else:
        key = name_or_path.split("/")[-1]
        if key not in KNOWN_MODELS:
            raise ValueError(
                f"unknown model {name_or_path!r}: not a local directory with config.json and not one of "
                f"{sorted(KNOWN_MODELS)} (no network: hub downloads are unavailable)")
        cfg = KNOWN_MODELS[key]()
        cfg.name_or_path = name_or_path
    if num_labels is not None:
        cfg.num_labels = num_labels
    return cfg
