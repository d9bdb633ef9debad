"""``save_pretrained``-layout checkpoints (SURVEY.md §2.1 C19, §3.5, §5 checkpoint/resume).

The reference ends with ``model.save_pretrained(dir); tokenizer.save_pretrained(dir)`` on
every rank (scripts/train.py:181-183).  Here rank 0 writes ``config.json`` +
``model.safetensors`` (HF PyTorch parameter names, fp32, metadata ``format=pt``) and the
tokenizer files when a tokenizer exists, and every rank then passes a barrier — the
reference's "all ranks write the same dir" race is not reproduced.  ``--model_name_or_path
<dir>`` reloads it, which is the reference's only form of resume; ``save_training_state``
adds optimizer state + step for a true resume.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Optional

import torch

from .models.configuration import EncoderConfig
from .models.hf_names import canonical_to_hf, hf_to_canonical

WEIGHTS_NAME = "model.safetensors"
TRAINING_STATE_NAME = "training_state.pt"


def save_pretrained(save_dir: str, cfg: EncoderConfig, canonical_state: Dict[str, torch.Tensor],
                    tokenizer=None) -> None:
    from safetensors.torch import save_file

    os.makedirs(save_dir, exist_ok=True)
    hf = {k: v.detach().to("cpu", torch.float32).contiguous() for k, v in canonical_to_hf(cfg, canonical_state).items()}
    tmp = os.path.join(save_dir, WEIGHTS_NAME + ".tmp")
    save_file(hf, tmp, metadata={"format": "pt"})
    os.replace(tmp, os.path.join(save_dir, WEIGHTS_NAME))
    with open(os.path.join(save_dir, "config.json"), "w") as f:
        json.dump(cfg.to_hf_dict(), f, indent=2, sort_keys=True)
        f.write("\n")
    if tokenizer is not None:
        tokenizer.save_pretrained(save_dir)


def load_pretrained_state(load_dir: str, cfg: EncoderConfig) -> Optional[Dict[str, torch.Tensor]]:
    """Canonical state dict from ``load_dir`` (safetensors or pytorch_model.bin), or None."""
    st = os.path.join(load_dir, WEIGHTS_NAME)
    pt = os.path.join(load_dir, "pytorch_model.bin")
    if os.path.exists(st):
        from safetensors.torch import load_file

        hf = load_file(st)
    elif os.path.exists(pt):
        hf = torch.load(pt, map_location="cpu", weights_only=True)
    else:
        return None
    return hf_to_canonical(cfg, hf)


def save_training_state(save_dir: str, optimizer_state: dict, step: int, epoch: int) -> None:
    os.makedirs(save_dir, exist_ok=True)
    torch.save({"optimizer": optimizer_state, "step": step, "epoch": epoch},
               os.path.join(save_dir, TRAINING_STATE_NAME))


def load_training_state(load_dir: str) -> Optional[dict]:
    p = os.path.join(load_dir, TRAINING_STATE_NAME)
    if not os.path.exists(p):
        return None
    # tensors, ints and dicts only: the restricted unpickler is enough and a checkpoint directory from an
    # untrusted source cannot run code through this path
    return torch.load(p, map_location="cpu", weights_only=True)
