"""``tf.distribute.MirroredStrategy``-shaped single-command front-end
(scripts/singe_node_train.py:40-41).

The reference starts ONE python process that drives every visible GPU.  Semantics kept:
the batch size given to ``fit`` is the *global* batch and is split across replicas, there is
no learning-rate scaling and no explicit broadcast step, gradients are combined by an
all-reduce every step.  With 0 or 1 device this is plain single-device training — on a
machine without GPUs it is the CPU oracle path of BASELINE config #1.

B200-native execution model: one process per GPU.  So with N > 1 visible GPUs
``MirroredStrategy.ensure_replicas()`` turns the single command into N replica processes
(same argv, ``RANK/LOCAL_RANK/WORLD_SIZE/MASTER_*`` set, stdout prefixed, whole job killed
if a replica dies) and the parent just waits — the user still types one command and sees one
job.  Inside a replica ``scope()`` joins the process group; the replicas then run the same
engine and the same fused NVLink all-reduce kernels as ``train.py``.
"""
from __future__ import annotations

import contextlib
import os
import sys
from typing import List, Optional

import torch

_CURRENT = None
CHILD_ENV = "HSTD_MIRRORED_REPLICA"


class MirroredStrategy:
    def __init__(self, devices: Optional[List[str]] = None):
        if devices is None:
            n = torch.cuda.device_count() if torch.cuda.is_available() and os.environ.get("HSTD_FORCE_CPU") != "1" else 0
            devices = [f"cuda:{i}" for i in range(n)] or ["cpu"]
        self.devices = [torch.device(d) for d in devices]
        self.is_replica_process = os.environ.get(CHILD_ENV) == "1"
        if self.is_replica_process:
            self.replica_id = int(os.environ.get("RANK", "0"))
            self._num = int(os.environ.get("WORLD_SIZE", "1"))
        else:
            self.replica_id, self._num = 0, len(self.devices)

    @property
    def num_replicas_in_sync(self) -> int:
        return self._num

    def per_replica_batch(self, global_batch: int) -> int:
        """Keras splits the global batch across replicas (scripts/singe_node_train.py:86-91)."""
        n = self.num_replicas_in_sync
        if global_batch % n:
            raise ValueError(f"global batch {global_batch} is not divisible by {n} replicas")
        return global_batch // n

    def ensure_replicas(self, argv: Optional[List[str]] = None, timeout_s: Optional[float] = None) -> bool:
        """With more than one device and not yet inside a replica: run this same command once
        per device and wait.  Returns True in the parent after the replicas finished (the
        caller should then exit), False when the caller *is* the (only) training process."""
        if self.is_replica_process or len(self.devices) <= 1:
            return False
        from ..launcher import _free_port, wait_all_or_kill, spawn_ranks

        argv = [sys.executable] + (argv if argv is not None else sys.argv)
        env = dict(os.environ)
        env[CHILD_ENV] = "1"
        procs, threads = spawn_ranks(argv, env, len(self.devices), _free_port())
        rc = wait_all_or_kill(procs, timeout_s)
        for t in threads:
            t.join(timeout=5)
        if rc != 0:
            raise SystemExit(rc)
        return True

    @contextlib.contextmanager
    def scope(self):
        global _CURRENT
        prev, _CURRENT = _CURRENT, self
        joined = False
        if self.is_replica_process and self._num > 1:
            from . import comm

            comm.init()
            joined = True
        try:
            yield self
        finally:
            _CURRENT = prev
            if joined:
                from . import comm

                comm.shutdown()


def current_strategy() -> "MirroredStrategy | None":
    return _CURRENT
