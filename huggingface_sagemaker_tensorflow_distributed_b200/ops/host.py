"""ctypes front-ends of the native *host* runtime in libhstd_b200.so: WordPiece tokenizer
(csrc/tokenizer.cpp), fusion-bucket manager (csrc/bucket_manager.cpp) and prefetching batch
loader (csrc/batch_loader.cpp).  None of these touch the GPU, so they load and run on a CPU-only
machine as long as the library has been built (``python -m <package>.build``).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _lib


@_lib.registrar
def _register(lib) -> None:
    vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
    lib.hstd_tok_create.argtypes = [C.c_char_p, i32, i32]
    lib.hstd_tok_create.restype = vp
    lib.hstd_tok_destroy.argtypes = [vp]
    lib.hstd_tok_destroy.restype = None
    lib.hstd_tok_vocab_size.argtypes = [vp]
    lib.hstd_tok_token_to_id.argtypes = [vp, C.c_char_p]
    lib.hstd_tok_id_to_token.argtypes = [vp, i32]
    lib.hstd_tok_id_to_token.restype = C.c_char_p
    lib.hstd_tok_special_ids.argtypes = [vp, vp]
    lib.hstd_tok_encode_batch.argtypes = [vp, C.POINTER(C.c_char_p), i32, i32, i32, vp, vp, vp, i32]

    lib.hstd_bpe_create.argtypes = [C.c_char_p, C.c_char_p]
    lib.hstd_bpe_create.restype = vp
    lib.hstd_bpe_destroy.argtypes = [vp]
    lib.hstd_bpe_destroy.restype = None
    lib.hstd_bpe_vocab_size.argtypes = [vp]
    lib.hstd_bpe_token_to_id.argtypes = [vp, C.c_char_p]
    lib.hstd_bpe_id_to_token.argtypes = [vp, i32]
    lib.hstd_bpe_id_to_token.restype = C.c_char_p
    lib.hstd_bpe_special_ids.argtypes = [vp, vp]
    lib.hstd_bpe_encode_batch.argtypes = [vp, C.POINTER(C.c_char_p), i32, i32, i32, vp, vp, vp, i32]

    lib.hstd_bucket_plan.argtypes = [vp, vp, i32, i64, vp, vp, vp, i32]
    lib.hstd_bucket_tracker_create.argtypes = [vp, vp, i32, i32]
    lib.hstd_bucket_tracker_create.restype = vp
    lib.hstd_bucket_tracker_destroy.argtypes = [vp]
    lib.hstd_bucket_tracker_destroy.restype = None
    lib.hstd_bucket_tracker_reset.argtypes = [vp]
    lib.hstd_bucket_tracker_reset.restype = None
    lib.hstd_bucket_tracker_mark.argtypes = [vp, vp, i32, vp]
    lib.hstd_bucket_tracker_done.argtypes = [vp]

    lib.hstd_loader_create.argtypes = [vp, vp, vp, i64, i32, vp, i64, i32, i32, i32, vp, vp, vp, C.c_int32, C.c_int32]
    lib.hstd_loader_create.restype = vp
    lib.hstd_loader_destroy.argtypes = [vp]
    lib.hstd_loader_destroy.restype = None
    lib.hstd_loader_start_epoch.argtypes = [vp]
    lib.hstd_loader_next.argtypes = [vp]
    lib.hstd_loader_release.argtypes = [vp, i32]


def _err(what: str) -> "_lib.HstdError":
    return _lib.HstdError(f"{what}: {_lib.load().hstd_last_error().decode(errors='replace')}")


# ------------------------------------------------------------------------------------------ tokenizer
class NativeWordPiece:
    """Owns one ``hstd_tokenizer``.  ``encode_batch`` returns dense int32 arrays."""

    def __init__(self, vocab_path: str, do_lower_case: bool = True, strip_accents: Optional[bool] = None):
        self._lib = _lib.load()
        strip = do_lower_case if strip_accents is None else strip_accents
        self._h = self._lib.hstd_tok_create(str(vocab_path).encode(), int(do_lower_case), int(strip))
        if not self._h:
            raise _err("hstd_tok_create")
        sp = (C.c_int32 * 4)()
        _lib.check(self._lib.hstd_tok_special_ids(self._h, sp), "hstd_tok_special_ids")
        self.pad_id, self.unk_id, self.cls_id, self.sep_id = (int(x) for x in sp)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._lib.hstd_tok_destroy(h)

    @property
    def vocab_size(self) -> int:
        return int(self._lib.hstd_tok_vocab_size(self._h))

    def token_to_id(self, token: str) -> int:
        return int(self._lib.hstd_tok_token_to_id(self._h, token.encode()))

    def id_to_token(self, idx: int) -> Optional[str]:
        s = self._lib.hstd_tok_id_to_token(self._h, int(idx))
        return None if s is None else s.decode(errors="replace")

    def encode_batch(self, texts: Sequence[str], max_len: int, add_special_tokens: bool = True,
                     num_threads: int = 0) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
        n = len(texts)
        ids = np.empty((n, max_len), dtype=np.int32)
        mask = np.empty((n, max_len), dtype=np.int32)
        lengths = np.empty((n,), dtype=np.int32)
        if n == 0:
            return ids, mask, lengths
        arr = (C.c_char_p * n)(*[t.replace("\x00", " ").encode("utf-8", errors="replace") for t in texts])
        rc = self._lib.hstd_tok_encode_batch(self._h, arr, n, int(max_len), int(add_special_tokens),
                                             ids.ctypes.data, mask.ctypes.data, lengths.ctypes.data, int(num_threads))
        _lib.check(rc, "hstd_tok_encode_batch")
        return ids, mask, lengths


class NativeBPE:
    """Owns one ``hstd_bpe`` (byte-level BPE, RoBERTa / GPT-2 files).  Same surface as NativeWordPiece;
    ``cls_id`` / ``sep_id`` are the <s> / </s> ids."""

    def __init__(self, vocab_json: str, merges_txt: str):
        self._lib = _lib.load()
        self._h = self._lib.hstd_bpe_create(str(vocab_json).encode(), str(merges_txt).encode())
        if not self._h:
            raise _err("hstd_bpe_create")
        sp = (C.c_int32 * 4)()
        _lib.check(self._lib.hstd_bpe_special_ids(self._h, sp), "hstd_bpe_special_ids")
        self.pad_id, self.unk_id, self.cls_id, self.sep_id = (int(x) for x in sp)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._lib.hstd_bpe_destroy(h)

    @property
    def vocab_size(self) -> int:
        return int(self._lib.hstd_bpe_vocab_size(self._h))

    def token_to_id(self, token: str) -> int:
        return int(self._lib.hstd_bpe_token_to_id(self._h, token.encode()))

    def id_to_token(self, idx: int) -> Optional[str]:
        s = self._lib.hstd_bpe_id_to_token(self._h, int(idx))
        return None if s is None else s.decode(errors="replace")

    def encode_batch(self, texts: Sequence[str], max_len: int, add_special_tokens: bool = True,
                     num_threads: int = 0) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
        n = len(texts)
        ids = np.empty((n, max_len), dtype=np.int32)
        mask = np.empty((n, max_len), dtype=np.int32)
        lengths = np.empty((n,), dtype=np.int32)
        if n == 0:
            return ids, mask, lengths
        arr = (C.c_char_p * n)(*[t.replace("\x00", " ").encode("utf-8", errors="replace") for t in texts])
        rc = self._lib.hstd_bpe_encode_batch(self._h, arr, n, int(max_len), int(add_special_tokens),
                                             ids.ctypes.data, mask.ctypes.data, lengths.ctypes.data, int(num_threads))
        _lib.check(rc, "hstd_bpe_encode_batch")
        return ids, mask, lengths


# ------------------------------------------------------------------------------------------ buckets
def bucket_plan(offsets: Sequence[int], ends: Sequence[int], bucket_elems: int) -> Tuple[List[Tuple[int, int]], List[int]]:
    """[(start, end)] per bucket and the bucket index of every parameter (arena order)."""
    lib = _lib.load()
    n = len(offsets)
    off = np.asarray(offsets, dtype=np.int64)
    end = np.asarray(ends, dtype=np.int64)
    starts = np.zeros(n, dtype=np.int64)
    bends = np.zeros(n, dtype=np.int64)
    pb = np.zeros(n, dtype=np.int32)
    nb = lib.hstd_bucket_plan(off.ctypes.data, end.ctypes.data, n, int(bucket_elems), starts.ctypes.data,
                              bends.ctypes.data, pb.ctypes.data, n)
    if nb < 0:
        raise _err("hstd_bucket_plan")
    return [(int(starts[i]), int(bends[i])) for i in range(nb)], [int(x) for x in pb]


class BucketTracker:
    def __init__(self, param_bucket: Sequence[int], exempt: Sequence[bool], num_buckets: int):
        self._lib = _lib.load()
        pb = np.asarray(param_bucket, dtype=np.int32)
        ex = np.asarray(exempt, dtype=np.uint8)
        self.num_buckets = int(num_buckets)
        self._h = self._lib.hstd_bucket_tracker_create(pb.ctypes.data, ex.ctypes.data, len(pb), self.num_buckets)
        if not self._h:
            raise _err("hstd_bucket_tracker_create")
        self._out = np.zeros(self.num_buckets, dtype=np.int32)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._lib.hstd_bucket_tracker_destroy(h)

    def reset(self) -> None:
        self._lib.hstd_bucket_tracker_reset(self._h)

    def mark(self, params: Sequence[int]) -> List[int]:
        p = np.asarray(list(params), dtype=np.int32)
        k = self._lib.hstd_bucket_tracker_mark(self._h, p.ctypes.data, len(p), self._out.ctypes.data)
        if k < 0:
            raise _err("hstd_bucket_tracker_mark")
        return [int(x) for x in self._out[:k]]

    @property
    def done(self) -> int:
        return int(self._lib.hstd_bucket_tracker_done(self._h))


# ------------------------------------------------------------------------------------------ loader
class NativeLoader:
    """Ring of caller-owned slots filled by a C++ worker thread (csrc/batch_loader.cpp).

    ``slots`` is a list of ``(ids, mask, labels)`` int32 tensors/arrays (pinned when a GPU is
    present); the tables must stay alive and unmoved for the lifetime of the loader."""

    def __init__(self, ids: np.ndarray, mask: np.ndarray, labels: np.ndarray, indices: np.ndarray, batch: int,
                 num_batches: int, slot_ptrs: Sequence[Tuple[int, int, int]], pad_id: int, ignore_label: int):
        self._lib = _lib.load()
        assert ids.dtype == np.int32 and mask.dtype == np.int32 and labels.dtype == np.int32
        assert ids.flags.c_contiguous and mask.flags.c_contiguous and labels.flags.c_contiguous
        self._keep = (ids, mask, labels)
        idx = np.ascontiguousarray(indices, dtype=np.int64)
        ns = len(slot_ptrs)
        pi = (C.c_void_p * ns)(*[p[0] for p in slot_ptrs])
        pm = (C.c_void_p * ns)(*[p[1] for p in slot_ptrs])
        pl = (C.c_void_p * ns)(*[p[2] for p in slot_ptrs])
        self._h = self._lib.hstd_loader_create(ids.ctypes.data, mask.ctypes.data, labels.ctypes.data, ids.shape[0],
                                               ids.shape[1], idx.ctypes.data, len(idx), int(batch), int(num_batches), ns,
                                               pi, pm, pl, int(pad_id), int(ignore_label))
        if not self._h:
            raise _err("hstd_loader_create")

    def close(self) -> None:
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._lib.hstd_loader_destroy(h)

    __del__ = close

    def start_epoch(self) -> None:
        _lib.check(self._lib.hstd_loader_start_epoch(self._h), "hstd_loader_start_epoch")

    def next(self) -> int:
        s = self._lib.hstd_loader_next(self._h)
        if s < -1:
            raise _err("hstd_loader_next")
        return int(s)

    def release(self, slot: int) -> None:
        _lib.check(self._lib.hstd_loader_release(self._h, int(slot)), "hstd_loader_release")
