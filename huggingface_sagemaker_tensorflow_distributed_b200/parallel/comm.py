"""Process-group plumbing: one process per GPU, ``torch.distributed`` (NCCL on GPUs, gloo on
CPU) for bootstrap / broadcast / barriers.  The gradient all-reduce hot path does not go
through here on GPUs — it is the hand-written NVLink kernel in csrc/allreduce_sm100.cu,
driven by ``parallel/dp_engine.py``; NCCL stays for bootstrap, the one-time parameter
broadcast and the baseline (SURVEY.md §2.3 N3/N4, §5 "Distributed communication backend").

Rank discovery accepts torchrun-style (RANK / WORLD_SIZE / LOCAL_RANK) and mpirun-style
(OMPI_COMM_WORLD_*) environments, mirroring how the reference is started either by plain
``python`` or ``mpirun`` (launch.py:19-24).  Without a launcher it degrades to size()==1,
as ``hvd.init()`` does.
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist


@dataclass
class CommState:
    rank: int = 0
    world_size: int = 1
    local_rank: int = 0
    local_size: int = 1
    backend: Optional[str] = None
    initialized: bool = False
    owns_process_group: bool = False


_STATE = CommState()


def _env_int(*names: str, default: int) -> int:
    for n in names:
        v = os.environ.get(n)
        if v is not None and v != "":
            return int(v)
    return default


def init(backend: Optional[str] = None, timeout_s: int = 600) -> CommState:
    """Idempotent.  Creates the default process group when WORLD_SIZE > 1."""
    global _STATE
    if _STATE.initialized:
        return _STATE
    rank = _env_int("RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", default=0)
    world = _env_int("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", default=1)
    local_rank = _env_int("LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", default=rank)
    local_size = _env_int("LOCAL_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_SIZE", default=world)
    use_cuda = torch.cuda.is_available() and os.environ.get("HSTD_FORCE_CPU", "0") != "1"
    if backend is None:
        backend = "nccl" if use_cuda else "gloo"
    if use_cuda:
        torch.cuda.set_device(local_rank % torch.cuda.device_count())  # GPU pinning (scripts/train.py:26-31)
    owns = False
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        kw = {}
        if backend == "nccl":
            kw["device_id"] = torch.device("cuda", local_rank % torch.cuda.device_count())
        dist.init_process_group(backend=backend, rank=rank, world_size=world,
                                timeout=datetime.timedelta(seconds=timeout_s), **kw)
        owns = True
    _STATE = CommState(rank, world, local_rank, local_size, backend if world > 1 else None, True, owns)
    return _STATE


def state() -> CommState:
    return _STATE


def shutdown() -> None:
    global _STATE
    if _STATE.owns_process_group and dist.is_initialized():
        dist.destroy_process_group()
    _STATE = CommState()


def rank() -> int:
    return _STATE.rank


def size() -> int:
    return _STATE.world_size


def local_rank() -> int:
    return _STATE.local_rank


def barrier() -> None:
    if _STATE.world_size > 1 and dist.is_initialized():
        if _STATE.backend == "nccl":
            dist.barrier(device_ids=[torch.cuda.current_device()])
        else:
            dist.barrier()


def broadcast_(tensor: torch.Tensor, root: int = 0) -> torch.Tensor:
    if _STATE.world_size > 1:
        dist.broadcast(tensor, src=root)
    return tensor


def allreduce_mean_(tensor: torch.Tensor) -> torch.Tensor:
    """Library all-reduce-average (baseline / CPU path / small metrics)."""
    if _STATE.world_size > 1:
        dist.all_reduce(tensor, op=dist.ReduceOp.SUM)
        tensor.div_(_STATE.world_size)
    return tensor


def allreduce_max_(tensor: torch.Tensor) -> torch.Tensor:
    if _STATE.world_size > 1:
        dist.all_reduce(tensor, op=dist.ReduceOp.MAX)
    return tensor


def device() -> torch.device:
    if torch.cuda.is_available() and os.environ.get("HSTD_FORCE_CPU", "0") != "1":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")
