"""Symmetric peer-mapped arenas + the fused all-reduce launcher (SURVEY.md §5
"Distributed communication backend", K9-K11).

Every rank allocates one device arena ``[weights bf16 | grads bf16 | flags u32]`` and maps the
arena of every peer:

  * preferred: ``torch.distributed._symmetric_memory`` rendezvous — gives the peer pointers
    *and*, when the fabric supports NVLS, a multicast address on which
    ``multimem.ld_reduce`` / ``multimem.st`` reduce / broadcast inside the NVSwitch;
  * fallback: our own CUDA-IPC exchange (``hstd_ipc_alloc`` / ``hstd_ipc_open``; handles
    travel through ``all_gather_object``) — no multicast, two-shot/one-shot only.

The kernels themselves (csrc/allreduce_sm100.cu) are ours either way; NCCL is not on this
path.  Per bucket the variant is chosen by size: one-shot below ``one_shot_bytes``, NVLS when
a multicast mapping exists, two-shot otherwise.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from ..ops import _lib
from . import comm

AR_LOCAL, AR_TWO_SHOT, AR_ONE_SHOT, AR_NVLS = 0, 1, 2, 3
MODE_NAMES = {AR_LOCAL: "local", AR_TWO_SHOT: "two_shot", AR_ONE_SHOT: "one_shot", AR_NVLS: "nvls"}


class AllreduceArgs(C.Structure):
    _fields_ = [("g_peers", C.c_void_p * 8), ("w_peers", C.c_void_p * 8), ("flag_peers", C.c_void_p * 8),
                ("g_mc", C.c_void_p), ("w_mc", C.c_void_p), ("epoch", C.c_void_p), ("error", C.c_void_p),
                ("master", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p), ("state", C.c_void_p),
                ("start", C.c_int64), ("numel", C.c_int64), ("world", C.c_int32), ("rank", C.c_int32),
                ("bucket_index", C.c_int32), ("mode", C.c_int32), ("update", C.c_int32), ("blocks", C.c_int32)]


@_lib.registrar
def _register(lib) -> None:
    lib.hstd_fused_allreduce.argtypes = [C.POINTER(AllreduceArgs), C.c_void_p]
    lib.hstd_fused_allreduce.restype = C.c_int
    lib.hstd_allreduce_max_blocks.restype = C.c_int
    lib.hstd_allreduce_flag_words.argtypes = [C.c_int]
    lib.hstd_allreduce_flag_words.restype = C.c_int64
    lib.hstd_ipc_alloc.argtypes = [C.c_int64, C.POINTER(C.c_void_p), C.c_void_p]
    lib.hstd_ipc_open.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
    lib.hstd_ipc_close.argtypes = [C.c_void_p]
    lib.hstd_ipc_free.argtypes = [C.c_void_p]
    lib.hstd_enable_peer_access.argtypes = [C.c_int, C.c_int]
    for n in ("hstd_ipc_alloc", "hstd_ipc_open", "hstd_ipc_close", "hstd_ipc_free", "hstd_enable_peer_access"):
        getattr(lib, n).restype = C.c_int


class _RawCudaBuffer:
    """Exposes a raw device pointer through __cuda_array_interface__ so torch can view it."""

    def __init__(self, ptr: int, nbytes: int):
        self.ptr, self.nbytes = ptr, nbytes
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


def _align(n: int, a: int) -> int:
    return (n + a - 1) // a * a


class SymmetricArena:
    """``nbytes`` of zeroed device memory on every rank, mapped into every rank."""

    def __init__(self, nbytes: int, device: torch.device, allow_symm_mem: bool = True):
        self.world, self.rank = comm.size(), comm.rank()
        self.nbytes = _align(nbytes, 2 << 20)
        self.device = device
        self.peer_ptrs: List[int] = []
        self.mc_ptr: int = 0
        self.kind = "none"
        self._keep = []
        err = None
        if allow_symm_mem and os.environ.get("HSTD_SYMM_MEM", "1") != "0":
            try:
                self._init_symm_mem()
            except Exception as e:  # noqa: BLE001 - any failure here means "use the IPC path"
                err = e
        if self.kind == "none":
            self._init_ipc()
            if err is not None and self.rank == 0 and os.environ.get("HSTD_DEBUG"):
                print(f"[p2p] symmetric-memory rendezvous unavailable ({type(err).__name__}: {err}); using CUDA IPC")

    def _init_symm_mem(self) -> None:
        import torch.distributed._symmetric_memory as symm_mem

        t = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=self.device)
        hdl = symm_mem.rendezvous(t, group=dist.group.WORLD.group_name)
        t.zero_()
        self.local = t
        self.peer_ptrs = [int(p) for p in hdl.buffer_ptrs]
        mc = 0
        try:
            mc = int(hdl.multicast_ptr or 0)
        except Exception:  # noqa: BLE001
            mc = 0
        if os.environ.get("HSTD_NVLS", "1") == "0":
            mc = 0
        self.mc_ptr = mc
        self._keep = [t, hdl]
        self.kind = "symm_mem"
        torch.cuda.synchronize()
        comm.barrier()

    def _init_ipc(self) -> None:
        lib = _lib.load()
        ptr = C.c_void_p()
        handle = (C.c_ubyte * 64)()
        _lib.check(lib.hstd_ipc_alloc(self.nbytes, C.byref(ptr), handle), "ipc_alloc")
        handles: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(handles, bytes(handle))
        self.peer_ptrs = []
        for r in range(self.world):
            if r == self.rank:
                self.peer_ptrs.append(int(ptr.value))
                continue
            h = (C.c_ubyte * 64).from_buffer_copy(handles[r])
            pp = C.c_void_p()
            _lib.check(lib.hstd_ipc_open(h, C.byref(pp)), "ipc_open")
            self.peer_ptrs.append(int(pp.value))
        self.local = torch.as_tensor(_RawCudaBuffer(int(ptr.value), self.nbytes), device=self.device)
        self.kind = "cuda_ipc"
        torch.cuda.synchronize()
        comm.barrier()

    def view(self, byte_offset: int, numel: int, dtype: torch.dtype) -> torch.Tensor:
        esz = torch.empty(0, dtype=dtype).element_size()
        return self.local[byte_offset:byte_offset + numel * esz].view(dtype)


class P2PGroup:
    def __init__(self, total_elems: int, buckets, device: torch.device, one_shot_bytes: int = 256 << 10,
                 blocks: int = 0, force_mode: Optional[int] = None):
        lib = _lib.load()
        self.world, self.rank, self.device = comm.size(), comm.rank(), device
        self.total, self.buckets = total_elems, buckets
        self.max_blocks = lib.hstd_allreduce_max_blocks()
        self.blocks = blocks or int(os.environ.get("HSTD_AR_BLOCKS", "48"))
        w_bytes = _align(total_elems * 2, 256)
        flag_words = lib.hstd_allreduce_flag_words(len(buckets) + 1)
        self.off_w, self.off_g, self.off_f = 0, w_bytes, 2 * w_bytes
        self.arena = SymmetricArena(2 * w_bytes + flag_words * 4, device)
        self.w16 = self.arena.view(self.off_w, total_elems, torch.bfloat16)
        self.g16 = self.arena.view(self.off_g, total_elems, torch.bfloat16)
        self.epochs = torch.zeros((len(buckets) + 1) * self.max_blocks, dtype=torch.int32, device=device)
        self.error = torch.zeros(1, dtype=torch.int32, device=device)
        env_mode = os.environ.get("HSTD_AR_MODE")
        if force_mode is None and env_mode:
            force_mode = {"two_shot": AR_TWO_SHOT, "one_shot": AR_ONE_SHOT, "nvls": AR_NVLS}[env_mode]
        self.modes: List[int] = []
        for b in buckets:
            if force_mode is not None:
                mode = force_mode
            elif b.numel * 2 <= one_shot_bytes:
                mode = AR_ONE_SHOT
            elif self.arena.mc_ptr:
                mode = AR_NVLS
            else:
                mode = AR_TWO_SHOT
            if mode == AR_NVLS and not self.arena.mc_ptr:
                mode = AR_TWO_SHOT
            self.modes.append(mode)
        # owned (fp32 master / moments) ranges, compact in bucket order
        self.owned: List[tuple] = []  # (arena_start, numel, compact_offset)
        off = 0
        for b, mode in zip(buckets, self.modes):
            if mode == AR_ONE_SHOT:
                self.owned.append((b.start, b.numel, off))
                off += b.numel
            else:
                n = b.numel // self.world
                self.owned.append((b.start + self.rank * n, n, off))
                off += n
        self.owned_elems = off
        self.comm_stream = torch.cuda.Stream(device=device)
        self._ev_pool: List[torch.cuda.Event] = []

    # ------------------------------------------------------------------ sharded state helpers
    def take_owned(self, flat: torch.Tensor) -> torch.Tensor:
        out = torch.empty(self.owned_elems, dtype=flat.dtype, device=flat.device)
        for (s, n, o) in self.owned:
            out[o:o + n] = flat[s:s + n]
        return out

    def gather_owned(self, compact: torch.Tensor) -> torch.Tensor:
        """Full [total] tensor on every rank from the per-rank owned pieces."""
        parts = [torch.empty_like(compact) for _ in range(self.world)]
        dist.all_gather(parts, compact.contiguous())
        full = torch.zeros(self.total, dtype=compact.dtype, device=compact.device)
        for b, mode, (_, n, o) in zip(self.buckets, self.modes, self.owned):
            if mode == AR_ONE_SHOT:
                full[b.start:b.start + n] = parts[0][o:o + n]
            else:
                for r in range(self.world):
                    full[b.start + r * n:b.start + (r + 1) * n] = parts[r][o:o + n]
        return full

    # ------------------------------------------------------------------ kernel launch
    def make_args(self, bucket_index: int, start: int, numel: int, mode: int, update: bool, state=None, master=None,
                  m=None, v=None, owned_offset: int = 0, blocks: Optional[int] = None) -> AllreduceArgs:
        a = AllreduceArgs()
        for r in range(self.world):
            base = self.arena.peer_ptrs[r]
            a.w_peers[r] = base + self.off_w
            a.g_peers[r] = base + self.off_g
            a.flag_peers[r] = base + self.off_f
        if self.arena.mc_ptr:
            a.w_mc = self.arena.mc_ptr + self.off_w
            a.g_mc = self.arena.mc_ptr + self.off_g
        a.epoch = self.epochs.data_ptr() + bucket_index * self.max_blocks * 4
        a.error = self.error.data_ptr()
        if update:
            a.master = master.data_ptr() + owned_offset * 4
            a.m = m.data_ptr() + owned_offset * 4
            a.v = v.data_ptr() + owned_offset * 4
            a.state = state.data_ptr()
        a.start, a.numel = start, numel
        a.world, a.rank, a.bucket_index, a.mode = self.world, self.rank, bucket_index, mode
        a.update, a.blocks = int(update), blocks or self.blocks
        return a

    def launch(self, args: AllreduceArgs, stream: Optional[torch.cuda.Stream] = None) -> None:
        _lib.check(_lib.load().hstd_fused_allreduce(C.byref(args), _lib.stream_ptr(stream)), "fused_allreduce")

    def check_errors(self) -> None:
        e = int(self.error.item())
        if e:
            raise RuntimeError(f"fused all-reduce watchdog: rank {self.rank} never heard from rank {e - 1} "
                               "(peer died or diverged); aborting instead of hanging")

    # ------------------------------------------------------------------ plan recording (called by DPEngine)
    def record_begin(self, plan) -> None:
        self._plan_events = []

    def record_bucket(self, plan, bucket, state, master, m, v, overlap: bool = True) -> None:
        mode = self.modes[bucket.index]
        start_o, n_o, off = self.owned[bucket.index]
        if getattr(self, "debug_no_comm", False):
            # timing aid for the "exposed communication" metric: the same optimizer work on the
            # owned slice, but no peer traffic and no cross-rank flags (ranks diverge: never train so)
            args = self.make_args(bucket.index, start_o, n_o, AR_LOCAL, True, state, master, m, v, off)
        else:
            args = self.make_args(bucket.index, bucket.start, bucket.numel, mode, True, state, master, m, v, off)
        plan.launches += 1
        if not overlap:
            plan.ops.append(lambda args=args: self.launch(args))
            plan.names.append(f"allreduce+adamw[{MODE_NAMES[mode]}]")
            return
        ev = torch.cuda.Event()
        self._plan_events.append(ev)
        cs = self.comm_stream

        def op(args=args, ev=ev):
            ev.record(torch.cuda.current_stream())
            cs.wait_event(ev)
            self.launch(args, cs)

        plan.ops.append(op)
        plan.names.append(f"allreduce+adamw[{MODE_NAMES[mode]}]@comm")
        plan.uses_comm_stream = True

    def record_end(self, plan) -> None:
        if not getattr(plan, "uses_comm_stream", False):
            return
        ev = torch.cuda.Event()
        self._plan_events.append(ev)
        cs = self.comm_stream

        def op(ev=ev):
            ev.record(cs)
            torch.cuda.current_stream().wait_event(ev)

        plan.ops.append(op)
        plan.names.append("join(comm)")
        plan.keep_events = self._plan_events

    def describe(self) -> Dict:
        return {"arena": self.arena.kind, "nvls": bool(self.arena.mc_ptr), "blocks": self.blocks,
                "buckets": [{"mb": round(b.numel * 2 / 2 ** 20, 2), "mode": MODE_NAMES[m]}
                            for b, m in zip(self.buckets, self.modes)]}
