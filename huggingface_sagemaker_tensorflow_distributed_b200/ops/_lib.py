"""ctypes binding of libhstd_b200.so (C ABI in include/hstd_b200.h).

The library is looked up in-tree (``<package>/lib/libhstd_b200.so``).  ``load()`` raises if
it is missing: on a GPU box the CUDA path is the only path, there is no silent fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_LIB = None
LIB_PATH = Path(__file__).resolve().parent.parent / "lib" / "libhstd_b200.so"


class HstdError(RuntimeError):
    pass


class GemmArgs(C.Structure):
    _fields_ = [
        ("A", C.c_void_p), ("B", C.c_void_p), ("C", C.c_void_p), ("bias", C.c_void_p), ("aux", C.c_void_p),
        ("lda", C.c_int64), ("a_bs0", C.c_int64), ("a_bs1", C.c_int64),
        ("ldb", C.c_int64), ("b_bs0", C.c_int64), ("b_bs1", C.c_int64),
        ("ldc", C.c_int64), ("c_bs0", C.c_int64), ("c_bs1", C.c_int64),
        ("ldaux", C.c_int64),
        ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32),
        ("nb0", C.c_int32), ("nb1", C.c_int32),
        ("a_mn_major", C.c_int32), ("b_mn_major", C.c_int32),
        ("epilogue", C.c_int32), ("alpha", C.c_float),
        ("block_n", C.c_int32), ("max_ctas", C.c_int32), ("cluster_m", C.c_int32),
    ]


def available() -> bool:
    return LIB_PATH.exists()


def load():
    """Returns the loaded library, building nothing.  Raises HstdError if the .so is absent."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = Path(os.environ.get("HSTD_B200_LIB", LIB_PATH))
    if not path.exists():
        raise HstdError(
            f"{path} not found — run `python -m huggingface_sagemaker_tensorflow_distributed_b200.build` "
            "(or __graft_entry__.build()) first; there is no non-CUDA fallback on a GPU box")
    lib = C.CDLL(str(path))
    lib.hstd_last_error.restype = C.c_char_p
    lib.hstd_version.restype = C.c_int
    lib.hstd_gemm_bf16.argtypes = [C.POINTER(GemmArgs), C.c_void_p]
    lib.hstd_gemm_bf16.restype = C.c_int
    lib.hstd_gemm_bf16_splitk.argtypes = [C.POINTER(GemmArgs), C.c_int, C.c_void_p, C.c_int64, C.c_void_p]
    lib.hstd_gemm_bf16_splitk.restype = C.c_int
    lib.hstd_gemm_splitk_workspace_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.hstd_gemm_splitk_workspace_bytes.restype = C.c_int64
    _bind_optional(lib)
    _LIB = lib
    return lib


def _bind_optional(lib) -> None:
    """Signatures of the remaining entry points; filled in by the ops modules' registrars."""
    for reg in _REGISTRARS:
        reg(lib)


_REGISTRARS: list = []


def registrar(fn):
    """Decorator: ``fn(lib)`` sets argtypes/restype for a family of entry points."""
    _REGISTRARS.append(fn)
    if _LIB is not None:
        fn(_LIB)
    return fn


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().hstd_last_error().decode(errors="replace")
        raise HstdError(f"{what or 'hstd call'} failed ({rc}): {msg}")


def stream_ptr(stream=None) -> int:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream
