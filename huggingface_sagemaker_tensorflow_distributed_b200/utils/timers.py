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


def kernel_timeline(step_fn, out_path, steps: int = 4) -> None:
    """Runs ``steps`` steps under torch.profiler (CUPTI activity records; works for CUDA-graph replays and for
    multi-rank jobs, where ncu cannot be used) and writes the kernels of the LAST step to ``out_path``:
    start offset, duration, stream, name — plus per-stream busy time and the step's span.  Profiling adds
    overhead; use the timeline for ordering / overlap / tails, not for absolute step time."""
    import torch
    from torch.profiler import ProfilerActivity, profile

    torch.cuda.synchronize()
    marks = []
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(steps):
            step_fn()
            torch.cuda.synchronize()
            marks.append(None)
    if out_path is None:
        return
    raw = prof.profiler.kineto_results.events()
    evs = [(e.start_ns() / 1e3, e.duration_ns() / 1e3, e.device_resource_id(), e.name()) for e in raw
           if e.device_type() == torch.autograd.DeviceType.CUDA]
    evs.sort()
    if not evs:
        with open(out_path, "w") as f:
            f.write("no CUDA activity records (CUPTI unavailable?)\n")
        return
    # split into steps at the largest gaps between consecutive kernels (the host synchronises between steps)
    gaps = sorted(range(1, len(evs)), key=lambda i: evs[i][0] - (evs[i - 1][0] + evs[i - 1][1]), reverse=True)
    cut = max(gaps[:steps - 1]) if steps > 1 and len(gaps) >= steps - 1 else 0
    last = evs[cut:]
    t0 = last[0][0]
    end = max(st + d for st, d, _, _ in last)
    streams = {}
    with open(out_path, "w") as f:
        f.write(f"# kernels of one step: start_us  dur_us  stream  name   (span {end - t0:.1f} us, {len(last)} records)\n")
        for st, d, sid, name in last:
            streams[sid] = streams.get(sid, 0.0) + d
            f.write(f"{st - t0:10.1f} {d:9.1f} {sid:>4d}  {name[:110]}\n")
        f.write("# busy us per stream: " + ", ".join(f"{k}: {v:.0f}" for k, v in sorted(streams.items())) + "\n")
