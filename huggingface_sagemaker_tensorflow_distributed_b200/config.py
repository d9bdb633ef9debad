"""Hyperparameter / environment contract (SURVEY.md §2.1 C3, C7 and layer B).

The reference turns an estimator's ``hyperparameters`` dict into ``--key value`` argv
(launch.py:13-18,51) which each script parses with argparse (scripts/train.py:36-52,
scripts/singe_node_train.py:13-29), with defaults for the directories taken from the
SageMaker container environment (``SM_OUTPUT_DATA_DIR``, ``SM_MODEL_DIR``, ``SM_NUM_GPUS``).

Differences from the reference, all deliberate:
  * ``--learning_rate`` is a float (the reference's ``type=str`` breaks ``lr * size``);
  * ``--do_train`` / ``--do_eval`` parse "false"/"0"/"no" as False (argparse ``type=bool``
    treats every non-empty string as True);
  * missing ``SM_*`` variables fall back to ``./output`` / ``./model`` instead of raising
    ``KeyError`` while the parser is built;
  * extra knobs the north star asks for: ``--weight_decay``, ``--lr_schedule``,
    ``--warmup_steps``, ``--max_seq_length``, ``--data_source``, ``--prefetch``, ``--no_shard``, ``--dtype``, ``--max_steps``,
    dataset sizing for the synthetic IMDB-shaped data, and bucket size for the all-reduce.
Unknown flags are ignored (``parse_known_args``), as in the reference.
"""
from __future__ import annotations

import argparse
import dataclasses
import json
import os
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Sequence

DEFAULT_OUTPUT_DIR = "./output"
DEFAULT_MODEL_DIR = "./model"


def str2bool(v: Any) -> bool:
    if isinstance(v, bool):
        return v
    s = str(v).strip().lower()
    if s in ("1", "true", "t", "yes", "y", "on"):
        return True
    if s in ("0", "false", "f", "no", "n", "off", ""):
        return False
    raise argparse.ArgumentTypeError(f"expected a boolean, got {v!r}")


@dataclass
class TrainConfig:
    # --- flags that exist in the reference (scripts/train.py:39-50) ---
    epochs: int = 3
    train_batch_size: int = 8
    eval_batch_size: int = 4
    model_name_or_path: Optional[str] = None
    learning_rate: float = 5e-5
    do_train: bool = True
    do_eval: bool = True
    output_data_dir: str = DEFAULT_OUTPUT_DIR
    model_dir: str = DEFAULT_MODEL_DIR
    n_gpus: Optional[str] = None  # parsed but unused, as in the reference (scripts/train.py:50)
    # --- additions ---
    weight_decay: float = 0.0
    lr_schedule: str = "constant"  # constant | linear
    warmup_steps: int = 0
    adam_epsilon: float = 1e-7  # Keras default
    max_seq_length: Optional[int] = None  # default: model max (512)
    no_shard: bool = False  # True = every rank iterates the full dataset (reference behaviour)
    dtype: str = "auto"  # auto: bf16 on CUDA, fp32 on CPU
    max_steps: int = -1  # stop each epoch after this many steps (steps_per_epoch, scripts/train.py:148)
    max_eval_steps: int = -1
    train_samples: int = 25000  # IMDB sizes (scripts/train.py:72)
    test_samples: int = 25000
    seed: int = 42
    dropout: Optional[float] = None  # override every dropout prob of the model (0 disables)
    bucket_mb: float = 32.0
    comm: str = "p2p"  # p2p (our NVLink kernels) | nccl (baseline plumbing)
    cuda_graph: bool = True
    log_every: int = 50
    metrics_jsonl: Optional[str] = None
    save_training_state: bool = False  # also write optimizer moments + step (true resume)
    # ids: synthetic token ids | text: synthetic reviews through the native tokenizer | hf:<name or dir>: load_dataset route
    data_source: str = "ids"
    prefetch: bool = True  # C++ batch loader thread + pinned ring (GPU runs)
    train_channel: Optional[str] = None  # SM_CHANNEL_TRAIN
    test_channel: Optional[str] = None  # SM_CHANNEL_TEST
    extra: Dict[str, Any] = field(default_factory=dict)

    def to_json(self) -> str:
        return json.dumps(dataclasses.asdict(self), sort_keys=True)


def build_parser(distributed: bool = True) -> argparse.ArgumentParser:
    """argparse parser with the reference's flag names.  ``distributed=False`` drops
    ``--n_gpus`` exactly as scripts/singe_node_train.py:27 does."""
    env = os.environ
    p = argparse.ArgumentParser()
    p.add_argument("--epochs", type=int, default=3)
    p.add_argument("--train_batch_size", type=int, default=8)
    p.add_argument("--eval_batch_size", type=int, default=4)
    p.add_argument("--model_name_or_path", type=str)
    p.add_argument("--learning_rate", type=float, default=5e-5)
    p.add_argument("--do_train", type=str2bool, default=True)
    p.add_argument("--do_eval", type=str2bool, default=True)
    p.add_argument("--output_data_dir", type=str, default=env.get("SM_OUTPUT_DATA_DIR", DEFAULT_OUTPUT_DIR))
    p.add_argument("--model_dir", type=str, default=env.get("SM_MODEL_DIR", DEFAULT_MODEL_DIR))
    if distributed:
        p.add_argument("--n_gpus", type=str, default=env.get("SM_NUM_GPUS"))
    p.add_argument("--weight_decay", type=float, default=0.0)
    p.add_argument("--lr_schedule", type=str, default="constant", choices=["constant", "linear"])
    p.add_argument("--warmup_steps", type=int, default=0)
    p.add_argument("--adam_epsilon", type=float, default=1e-7)
    p.add_argument("--max_seq_length", type=int, default=None)
    p.add_argument("--no_shard", type=str2bool, nargs="?", const=True, default=False)
    p.add_argument("--dtype", type=str, default="auto", choices=["auto", "bf16", "fp32"])
    p.add_argument("--max_steps", type=int, default=-1)
    p.add_argument("--max_eval_steps", type=int, default=-1)
    p.add_argument("--train_samples", type=int, default=25000)
    p.add_argument("--test_samples", type=int, default=25000)
    p.add_argument("--seed", type=int, default=42)
    p.add_argument("--dropout", type=float, default=None)
    p.add_argument("--bucket_mb", type=float, default=32.0)
    p.add_argument("--comm", type=str, default="p2p", choices=["p2p", "nccl"])
    p.add_argument("--cuda_graph", type=str2bool, default=True)
    p.add_argument("--log_every", type=int, default=50)
    p.add_argument("--metrics_jsonl", type=str, default=None)
    p.add_argument("--save_training_state", type=str2bool, nargs="?", const=True, default=False)
    p.add_argument("--data_source", type=str, default="ids",
                   help="ids | text | hf:<dataset name in the local HF cache or save_to_disk directory> (e.g. hf:imdb)")
    p.add_argument("--prefetch", type=str2bool, nargs="?", const=True, default=True)
    p.add_argument("--train_channel", type=str, default=env.get("SM_CHANNEL_TRAIN"))
    p.add_argument("--test_channel", type=str, default=env.get("SM_CHANNEL_TEST"))
    return p


def parse_args(argv: Optional[Sequence[str]] = None, distributed: bool = True) -> TrainConfig:
    ns, unknown = build_parser(distributed).parse_known_args(argv)
    d = vars(ns)
    cfg = TrainConfig(**{k: v for k, v in d.items() if k in TrainConfig.__dataclass_fields__})
    cfg.extra = {"unknown_args": list(unknown)}
    return cfg


# --------------------------------------------------------------------------- launcher side
def hyperparameters_to_argv(hyperparameters: Dict[str, Any]) -> List[str]:
    """SageMaker's dict → ``--key value`` transformation (launch.py:13-18,51)."""
    argv: List[str] = []
    for k, v in hyperparameters.items():
        argv.append(f"--{k}")
        argv.append(json.dumps(v) if isinstance(v, (dict, list)) else str(v))
    return argv


def sagemaker_env(*, model_dir: str, output_data_dir: str, num_gpus: int, hyperparameters: Dict[str, Any],
                  channels: Optional[Dict[str, str]] = None, smddp: bool = False, hosts: Sequence[str] = ("algo-1",),
                  current_host: str = "algo-1", job_name: str = "local-job") -> Dict[str, str]:
    """Environment a SageMaker training container would provide (SURVEY.md §1 layer B)."""
    env = {
        "SM_MODEL_DIR": model_dir,
        "SM_OUTPUT_DATA_DIR": output_data_dir,
        "SM_OUTPUT_DIR": os.path.dirname(output_data_dir.rstrip("/")) or output_data_dir,
        "SM_NUM_GPUS": str(num_gpus),
        "SM_NUM_CPUS": str(os.cpu_count() or 1),
        "SM_HPS": json.dumps(hyperparameters),
        "SM_HOSTS": json.dumps(list(hosts)),
        "SM_CURRENT_HOST": current_host,
        "SM_CHANNELS": json.dumps(sorted((channels or {}).keys())),
        "SM_FRAMEWORK_PARAMS": json.dumps({"sagemaker_distributed_dataparallel_enabled": True} if smddp else {}),
        "TRAINING_JOB_NAME": job_name,
    }
    for name, path in (channels or {}).items():
        env[f"SM_CHANNEL_{name.upper()}"] = path
    for k, v in hyperparameters.items():
        env[f"SM_HP_{k.upper()}"] = str(v)
    return env


def is_sagemaker_dp_enabled(env: Optional[Dict[str, str]] = None) -> bool:
    """Same test the reference imports from transformers (scripts/train.py:13,16): the smddp
    flag inside the ``SM_FRAMEWORK_PARAMS`` JSON."""
    env = os.environ if env is None else env
    try:
        params = json.loads(env.get("SM_FRAMEWORK_PARAMS", "{}"))
    except json.JSONDecodeError:
        return False
    return bool(params.get("sagemaker_distributed_dataparallel_enabled", False))
