"""Device-side timing and tracing helpers (SURVEY.md §5 "Tracing / profiling": the reference
has only a wall clock around ``fit``, scripts/train.py:142,154).

* ``CudaStepTimer``  — CUDA-event timing of a region on the launching stream, reduced with
  max over ranks (multi-GPU numbers are never wall-clock numbers).
* ``nvtx_range``     — NVTX ranges (no-ops without CUDA) so an Nsight timeline is readable.
* ``profile_ops``    — per-op device time of an engine plan from event pairs around every
  recorded kernel call; what ``tools/op_profile.py`` prints.
"""
from __future__ import annotations

import contextlib
from collections import defaultdict
from typing import Callable, Dict, List, Tuple

import torch


class CudaStepTimer:
    def __init__(self):
        self.e0 = torch.cuda.Event(enable_timing=True)
        self.e1 = torch.cuda.Event(enable_timing=True)

    def __enter__(self):
        from ..parallel import comm
        comm.barrier()
        torch.cuda.synchronize()
        self.e0.record()
        return self

    def __exit__(self, *exc):
        from ..parallel import comm
        self.e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([self.e0.elapsed_time(self.e1)], dtype=torch.float64, device="cuda")
        comm.allreduce_max_(t)
        self.ms = t.item()
        return False


@contextlib.contextmanager
def nvtx_range(name: str):
    if torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


def profile_ops(ops: List[Callable[[], None]], names: List[str], repeats: int = 5) -> List[Tuple[str, int, float]]:
    """[(name, calls per step, total us per step)] sorted by time; eager execution, one event
    pair per op, a single synchronize per repeat."""
    acc: Dict[str, List[float]] = defaultdict(lambda: [0, 0.0])
    for _ in range(2):
        for f in ops:
            f()
    torch.cuda.synchronize()
    for _ in range(repeats):
        evs = []
        for f in ops:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            f()
            b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
        for n, (a, b) in zip(names, evs):
            acc[n][0] += 1
            acc[n][1] += a.elapsed_time(b) * 1e3
    out = [(n, c // repeats, t / repeats) for n, (c, t) in acc.items()]
    out.sort(key=lambda x: -x[2])
    return out
