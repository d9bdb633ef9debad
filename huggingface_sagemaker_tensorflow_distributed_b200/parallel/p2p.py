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
                ("bucket_index", C.c_int32), ("mode", C.c_int32), ("update", C.c_int32), ("blocks", C.c_int32),
                ("row_stamp", C.c_void_p), ("ids_peers", C.c_void_p * 8), ("row_elems", C.c_int32),
                ("rows", C.c_int32), ("n_ids", C.c_int32), ("dense_rows", C.c_int32), ("vocab_rows", C.c_int32),
                ("tail", C.c_int32), ("row_elem0", C.c_int64),
                ("phase", C.c_int32), ("end_barrier", C.c_int32)]


ROWS_ALL, ROWS_LIVE, ROWS_IDLE, ROWS_BOTH = 0, 1, 2, 3
MAX_IDS_PER_RANK = 64 * 1024  # tokens of one step and rank the peer-visible id slot can hold


@_lib.registrar
def _register(lib) -> None:
    lib.hstd_fused_allreduce.argtypes = [C.POINTER(AllreduceArgs), C.c_void_p]
    lib.hstd_fused_allreduce.restype = C.c_int
    lib.hstd_sparse_rows_mark.argtypes = [C.POINTER(AllreduceArgs), C.c_void_p]
    lib.hstd_sparse_rows_mark.restype = C.c_int
    for n in ("hstd_fused_allreduce_emulated", "hstd_sparse_rows_mark_emulated"):
        getattr(lib, n).argtypes = [C.POINTER(AllreduceArgs), C.c_int, C.c_void_p]
        getattr(lib, n).restype = C.c_int
    lib.hstd_allreduce_max_blocks.restype = C.c_int
    lib.hstd_allreduce_flag_words.argtypes = [C.c_int]
    lib.hstd_allreduce_flag_words.restype = C.c_int64
    lib.hstd_ipc_alloc.argtypes = [C.c_int64, C.POINTER(C.c_void_p), C.c_void_p]
    lib.hstd_ipc_open.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
    lib.hstd_ipc_close.argtypes = [C.c_void_p]
    lib.hstd_ipc_free.argtypes = [C.c_void_p]
    for n in ("hstd_ipc_alloc", "hstd_ipc_open", "hstd_ipc_close", "hstd_ipc_free"):
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
        # Blocks (of 128 threads, small enough to sit next to a GEMM CTA on the same SM) per bucket kernel.
        # The trailing bucket(s) have no compute left to protect and take the whole budget.
        self.blocks = blocks or int(os.environ.get("HSTD_AR_BLOCKS", "128"))
        self.tail_blocks = int(os.environ.get("HSTD_AR_TAIL_BLOCKS", str(self.max_blocks)))
        # split schedule: reduce underneath backward, AdamW of all buckets at full HBM speed at the end of the step
        self.split = os.environ.get("HSTD_AR_SPLIT", "1") != "0"
        w_bytes = _align(total_elems * 2, 256)
        # flag slots: one per bucket, one spare (collective tests / sweeps), one for the sparse-row marking
        self.mark_slot = len(buckets) + 1
        flag_words = lib.hstd_allreduce_flag_words(len(buckets) + 2)
        self.off_w, self.off_g, self.off_f = 0, w_bytes, 2 * w_bytes
        self.off_ids = _align(self.off_f + flag_words * 4, 256)
        self.arena = SymmetricArena(self.off_ids + MAX_IDS_PER_RANK * 4, device)
        self.w16 = self.arena.view(self.off_w, total_elems, torch.bfloat16)
        self.g16 = self.arena.view(self.off_g, total_elems, torch.bfloat16)
        self.epochs = torch.zeros((len(buckets) + 2) * self.max_blocks, dtype=torch.int32, device=device)
        self.sparse = next((b for b in buckets if getattr(b, "sparse", False)), None)
        self.row_stamp = (torch.zeros(self.sparse.numel // self.sparse.row_elems, dtype=torch.int32, device=device)
                          if self.sparse is not None else None)
        self.error = torch.zeros(1, dtype=torch.int32, device=device)
        self._poll_host = None
        env_mode = os.environ.get("HSTD_AR_MODE")
        if force_mode is None and env_mode:
            force_mode = {"two_shot": AR_TWO_SHOT, "one_shot": AR_ONE_SHOT, "nvls": AR_NVLS}[env_mode]
        self.modes: List[int] = []
        # Measured (profiles/scale_r2_n2.md): with two ranks the in-switch reduction has nothing to save — one
        # peer load per vector either way — and plain peer loads / stores were faster (2.417 vs 2.436 ms/step).
        nvls_pays = bool(self.arena.mc_ptr) and self.world > 2
        for b in buckets:
            if force_mode is not None:
                mode = force_mode
            elif getattr(b, "sparse", False):  # sharded state is what the row filters work on
                mode = AR_NVLS if nvls_pays else AR_TWO_SHOT
            elif b.numel * 2 <= one_shot_bytes:
                mode = AR_ONE_SHOT
            elif nvls_pays:
                mode = AR_NVLS
            else:
                mode = AR_TWO_SHOT
            if mode == AR_NVLS and not self.arena.mc_ptr:
                mode = AR_TWO_SHOT
            if mode == AR_ONE_SHOT and getattr(b, "sparse", False):
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
    def ids_view(self, n: int) -> torch.Tensor:
        """int32 [n] view of this rank's peer-visible token-id slot: the step's input_ids are copied straight
        into it, so the sparse-row marking kernels of the other ranks can read them."""
        if n > MAX_IDS_PER_RANK:
            raise ValueError(f"batch of {n} tokens per rank exceeds the peer-visible id slot ({MAX_IDS_PER_RANK})")
        return self.arena.view(self.off_ids, n, torch.int32)

    def make_args(self, bucket_index: int, start: int, numel: int, mode: int, update: bool, state=None, master=None,
                  m=None, v=None, owned_offset: int = 0, blocks: Optional[int] = None, rows: int = ROWS_ALL,
                  world: Optional[int] = None, row_elem0: int = 0, tail: bool = False, phase: int = 0,
                  end_barrier: bool = False) -> AllreduceArgs:
        a = AllreduceArgs()
        a.tail, a.phase, a.end_barrier = int(tail), phase, int(end_barrier)
        if rows != ROWS_ALL or bucket_index == self.mark_slot:
            sp = self.sparse
            a.row_stamp = self.row_stamp.data_ptr()
            a.row_elems, a.rows, a.row_elem0 = sp.row_elems, rows, row_elem0
            a.dense_rows, a.vocab_rows = sp.dense_rows, sp.vocab_rows
            for r in range(self.world):
                a.ids_peers[r] = self.arena.peer_ptrs[r] + self.off_ids
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
        if state is not None:
            a.state = state.data_ptr()
        if update:
            a.master = master.data_ptr() + owned_offset * 4
            a.m = m.data_ptr() + owned_offset * 4
            a.v = v.data_ptr() + owned_offset * 4
        a.start, a.numel = start, numel
        a.world, a.rank, a.bucket_index, a.mode = self.world, self.rank, bucket_index, mode
        a.update, a.blocks = int(update), blocks or self.blocks
        if world == 1:
            # this rank on its own (exposed-communication timing): "peer 0" is the local arena
            a.world, a.rank = 1, 0
            base = self.arena.peer_ptrs[self.rank]
            a.w_peers[0], a.g_peers[0], a.flag_peers[0] = base + self.off_w, base + self.off_g, base + self.off_f
            a.ids_peers[0] = base + self.off_ids
        return a

    def launch_mark(self, args: AllreduceArgs, stream: Optional[torch.cuda.Stream] = None) -> None:
        _lib.check(_lib.load().hstd_sparse_rows_mark(C.byref(args), _lib.stream_ptr(stream)), "sparse_rows_mark")

    def launch(self, args: AllreduceArgs, stream: Optional[torch.cuda.Stream] = None) -> None:
        _lib.check(_lib.load().hstd_fused_allreduce(C.byref(args), _lib.stream_ptr(stream)), "fused_allreduce")

    def _raise(self, e: int) -> None:
        raise RuntimeError(f"fused all-reduce watchdog: rank {self.rank} never heard from rank {e - 1} "
                           "(peer died or diverged); aborting instead of hanging")

    def check_errors(self) -> None:
        """Blocking read of the watchdog word (end of fit / epoch)."""
        e = int(self.error.item())
        if e:
            self._raise(e)

    def poll_errors(self) -> None:
        """Non-blocking health check for the training loop: looks at the copy of the watchdog word that
        the PREVIOUS poll requested (pinned host memory, no stream synchronisation) and requests a new
        one.  A dead peer therefore surfaces within two polling intervals instead of at the end of fit."""
        if self._poll_host is None:
            self._poll_host = torch.zeros(1, dtype=torch.int32).pin_memory()
            self._poll_event = None
        if self._poll_event is not None:
            if not self._poll_event.query():
                return  # previous copy still in flight (the device may be stuck in a watchdog spin)
            e = int(self._poll_host.item())
            if e:
                self._raise(e)
        self._poll_host.copy_(self.error, non_blocking=True)
        self._poll_event = torch.cuda.Event()
        self._poll_event.record()

    # ------------------------------------------------------------------ plan recording (called by DPEngine)
    def record_begin(self, plan) -> None:
        pass

    def _fork_to_comm(self, plan, side_stream=None):
        """Returns a callable that makes the comm stream wait for everything recorded so far on the main
        stream and, when weight-gradient GEMMs run on the engine's side stream, on that stream too — the
        main stream itself is never held up for a bucket."""
        ev = torch.cuda.Event()
        plan.keep.append(ev)
        cs = self.comm_stream
        ev2 = None
        if side_stream is not None:
            ev2 = torch.cuda.Event()
            plan.keep.append(ev2)

        def fork():
            ev.record(torch.cuda.current_stream())
            cs.wait_event(ev)
            if ev2 is not None:
                ev2.record(side_stream)
                cs.wait_event(ev2)

        return fork

    def record_step_start(self, plan, state, master, m, v, n_ids: int) -> None:
        """Row-sparse embedding bucket, start of the step (called right after the step prologue), on the comm
        stream underneath the forward pass: mark the rows that are live in this step and clear them in the
        local gradient arena.  One-launch schedule (HSTD_AR_SPLIT=0) only: the idle rows also take their
        zero-gradient Adam step here."""
        b = self.sparse
        if b is None:
            return
        local = getattr(self, "debug_no_comm", False)
        mode = self.modes[b.index]
        start_o, n_o, off = self.owned[b.index]
        mark = self.make_args(self.mark_slot, b.start, b.numel, mode, False, state, blocks=32,
                              world=1 if local else None)
        mark.n_ids = n_ids
        idle = None
        if not self.split:
            if local:
                idle = self.make_args(b.index, start_o, n_o, AR_TWO_SHOT, True, state, master, m, v, off,
                                      rows=ROWS_IDLE, world=1, row_elem0=start_o - b.start)
            else:
                idle = self.make_args(b.index, b.start, b.numel, mode, True, state, master, m, v, off, rows=ROWS_IDLE)
        fork = self._fork_to_comm(plan)
        ev_marked = torch.cuda.Event()
        plan.keep.append(ev_marked)
        cs = self.comm_stream

        def op():
            fork()
            self.launch_mark(mark, cs)
            ev_marked.record(cs)
            if idle is not None:
                self.launch(idle, cs)

        plan.launches += 1 + (idle is not None)
        plan.ops.append(op)
        plan.names.append("sparse rows: mark" + (" + idle-row adamw" if idle is not None else "") + "@comm")
        plan.uses_comm_stream = True
        plan.rows_marked_event = ev_marked

    def record_wait_rows_marked(self, plan) -> None:
        """Main stream: the live gradient rows have been cleared (needed before the embedding scatter-add)."""
        ev = getattr(plan, "rows_marked_event", None)
        if ev is None:
            return
        plan.ops.append(lambda: torch.cuda.current_stream().wait_event(ev))
        plan.names.append("wait(rows marked)")

    def record_bucket(self, plan, bucket, state, master, m, v, overlap: bool = True, side_stream=None,
                      last: bool = False) -> None:
        """Called when the gradients of ``bucket`` are complete.  Split schedule (default): a reduce-only launch
        that runs underneath the rest of backward; the optimizer pass of every bucket follows at the end of the
        step (record_end).  One-launch schedule: reduce + AdamW + weight broadcast here."""
        mode = self.modes[bucket.index]
        start_o, n_o, off = self.owned[bucket.index]
        sparse = getattr(bucket, "sparse", False)
        rows = ROWS_LIVE if sparse else ROWS_ALL
        tail = last or sparse
        blocks = self.tail_blocks if tail else self.blocks
        local = getattr(self, "debug_no_comm", False)
        split = self.split and mode != AR_ONE_SHOT
        if split:
            if not hasattr(plan, "pending_updates"):
                plan.pending_updates = []
            plan.pending_updates.append(bucket)
            if local:
                return  # exposed-communication timing: no reduce at all, the update pass reads the local gradients
            args = self.make_args(bucket.index, bucket.start, bucket.numel, mode, True, state, master, m, v, off, blocks,
                                  rows=rows, tail=tail, phase=1)
            label = f"reduce[{MODE_NAMES[mode]}{',live rows' if sparse else ''}]"
        elif local:
            # timing aid for the "exposed communication" metric: the same optimizer work on the
            # owned slice, but no peer traffic and no cross-rank flags (ranks diverge: never train so)
            if sparse:
                args = self.make_args(bucket.index, start_o, n_o, AR_TWO_SHOT, True, state, master, m, v, off, blocks,
                                      rows=ROWS_LIVE, world=1, row_elem0=start_o - bucket.start, tail=tail)
            else:
                args = self.make_args(bucket.index, start_o, n_o, AR_LOCAL, True, state, master, m, v, off, blocks)
            label = "adamw[local]"
        else:
            args = self.make_args(bucket.index, bucket.start, bucket.numel, mode, True, state, master, m, v, off, blocks,
                                  rows=rows, tail=tail)
            label = f"allreduce+adamw[{MODE_NAMES[mode]}{',live rows' if sparse else ''}]"
        plan.launches += 1
        if not overlap:
            plan.ops.append(lambda args=args: self.launch(args))
            plan.names.append(label)
            return
        fork = self._fork_to_comm(plan, side_stream)
        cs = self.comm_stream

        def op(args=args, fork=fork):
            fork()
            self.launch(args, cs)

        plan.ops.append(op)
        plan.names.append(label + "@comm")
        plan.uses_comm_stream = True

    def record_end(self, plan, state=None, master=None, m=None, v=None, overlap: bool = True) -> None:
        """End of backward.  Split schedule: the optimizer pass of every sharded bucket, big blocks, back to back
        behind the last reduce (whose barrier told us that every rank has finished backward, so peers' weights
        may be overwritten); the last launch carries the step's closing barrier.  Then the main stream joins."""
        pending = getattr(plan, "pending_updates", [])
        local = getattr(self, "debug_no_comm", False)
        cs = self.comm_stream if (overlap and getattr(plan, "uses_comm_stream", False)) else None
        launches = []
        for k, bucket in enumerate(pending):
            mode = self.modes[bucket.index]
            start_o, n_o, off = self.owned[bucket.index]
            sparse = getattr(bucket, "sparse", False)
            rows = ROWS_BOTH if sparse else ROWS_ALL
            if local:
                args = self.make_args(bucket.index, start_o, n_o, AR_TWO_SHOT, True, state, master, m, v, off,
                                      self.tail_blocks, rows=rows, world=1, row_elem0=start_o - bucket.start, tail=True,
                                      phase=2)
            else:
                args = self.make_args(bucket.index, bucket.start, bucket.numel, mode, True, state, master, m, v, off,
                                      self.tail_blocks, rows=rows, tail=True, phase=2,
                                      end_barrier=(k == len(pending) - 1))
            launches.append(args)
        if launches:
            fork = self._fork_to_comm(plan) if (cs is not None and local) else None  # (no reduce launched a fork)

            def run(launches=launches, fork=fork):
                if fork is not None:
                    fork()
                for a in launches:
                    self.launch(a, cs)

            plan.launches += len(launches)
            plan.ops.append(run)
            plan.names.append(f"adamw + weight broadcast x{len(launches)}" + ("@comm" if cs is not None else ""))
        if not getattr(plan, "uses_comm_stream", False):
            return
        ev = torch.cuda.Event()
        plan.keep.append(ev)
        cs2 = self.comm_stream

        def op(ev=ev):
            ev.record(cs2)
            torch.cuda.current_stream().wait_event(ev)

        plan.ops.append(op)
        plan.names.append("join(comm)")

    def describe(self) -> Dict:
        return {"arena": self.arena.kind, "nvls": bool(self.arena.mc_ptr), "blocks": self.blocks,
                "tail_blocks": self.tail_blocks,
                "schedule": "reduce under backward, AdamW at step end" if self.split else "reduce+AdamW per bucket",
                "buckets": [{"mb": round(b.numel * 2 / 2 ** 20, 2), "mode": MODE_NAMES[m],
                             **({"row_sparse": True} if getattr(b, "sparse", False) else {})}
                            for b, m in zip(self.buckets, self.modes)]}


class EmulatedP2P:
    """``world`` ranks' arenas on ONE GPU, driven through the emulated entry points: a single cooperative
    grid runs every rank's blocks with the same per-block code as the real kernels (flag barriers, peer
    loads / stores, fused AdamW, row filters).  For tests and profiling on a box with fewer GPUs than ranks;
    separate launches that wait on one another are not an option on one device (nothing guarantees they are
    co-scheduled), hence the one-kernel form.  Two-shot and one-shot only (NVLS needs a multicast mapping)."""

    def __init__(self, world: int, total_elems: int, num_buckets: int, device: torch.device, row_elems: int = 0,
                 rows: int = 0, blocks: int = 8):
        lib = _lib.load()
        self.world, self.total, self.device, self.blocks = world, total_elems, device, blocks
        self.max_blocks = lib.hstd_allreduce_max_blocks()
        nflag = lib.hstd_allreduce_flag_words(num_buckets + 2)
        self.mark_slot = num_buckets + 1
        z = lambda n, dt: [torch.zeros(n, dtype=dt, device=device) for _ in range(world)]  # noqa: E731
        self.w16, self.g16 = z(total_elems, torch.bfloat16), z(total_elems, torch.bfloat16)
        self.flags = z(nflag, torch.int32)
        self.epochs = z((num_buckets + 2) * self.max_blocks, torch.int32)
        self.errors = z(1, torch.int32)
        self.ids = z(MAX_IDS_PER_RANK, torch.int32)
        self.row_elems = row_elems
        self.stamps = z(max(rows, 1), torch.int32)

    def args(self, rank: int, bucket_index: int, start: int, numel: int, mode: int, update: bool, state=None,
             master=None, m=None, v=None, rows: int = ROWS_ALL, n_ids: int = 0, dense_rows: int = 0,
             vocab_rows: int = 0, phase: int = 0, end_barrier: bool = False) -> AllreduceArgs:
        a = AllreduceArgs()
        a.phase, a.end_barrier = phase, int(end_barrier)
        for r in range(self.world):
            a.w_peers[r], a.g_peers[r] = self.w16[r].data_ptr(), self.g16[r].data_ptr()
            a.flag_peers[r], a.ids_peers[r] = self.flags[r].data_ptr(), self.ids[r].data_ptr()
        a.epoch = self.epochs[rank].data_ptr() + bucket_index * self.max_blocks * 4
        a.error = self.errors[rank].data_ptr()
        if state is not None:
            a.state = state.data_ptr()
        if update:
            a.master, a.m, a.v = master.data_ptr(), m.data_ptr(), v.data_ptr()
        a.start, a.numel, a.world, a.rank = start, numel, self.world, rank
        a.bucket_index, a.mode, a.update, a.blocks = bucket_index, mode, int(update), self.blocks
        # rank 0's stamps serve every emulated rank in the all-reduce (the marking leaves them identical)
        a.row_stamp = self.stamps[rank if bucket_index == self.mark_slot else 0].data_ptr()
        a.row_elems, a.rows, a.n_ids, a.dense_rows, a.vocab_rows = self.row_elems, rows, n_ids, dense_rows, vocab_rows
        return a

    def _array(self, per_rank):
        arr = (AllreduceArgs * self.world)()
        for r, a in enumerate(per_rank):
            arr[r] = a
        return arr

    def allreduce(self, per_rank) -> None:
        _lib.check(_lib.load().hstd_fused_allreduce_emulated(self._array(per_rank), self.world, _lib.stream_ptr(None)),
                   "fused_allreduce_emulated")

    def mark(self, per_rank) -> None:
        _lib.check(_lib.load().hstd_sparse_rows_mark_emulated(self._array(per_rank), self.world, _lib.stream_ptr(None)),
                   "sparse_rows_mark_emulated")

    def check_errors(self) -> None:
        for r, e in enumerate(self.errors):
            assert int(e.item()) == 0, f"emulated rank {r}: watchdog fired ({int(e.item())})"
