"""hstd_colsum_noalloc: the column-sum entry point for callers that cannot provide scratch
memory and must not trigger a device allocation (a bare C caller without an allocator, or a
process whose GPU memory already belongs to some other pool).  It is called here the way
such a caller would — straight through the C ABI with raw pointers — and must give what the
workspace-taking hstd_colsum gives, also when called twice in a row with different widths
(the built-in scratch still holds the first call's partial sums during the second).

Collected last on purpose: a run that stops at the first failure has then already run
everything else.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import _lib as L

    lib = L.load()
    lib.hstd_colsum_noalloc.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    lib.hstd_colsum_noalloc.restype = C.c_int
    lib.hstd_colsum_noalloc_max_n.restype = C.c_int
    lib.hstd_colsum.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.hstd_colsum.restype = C.c_int
    lib.hstd_colsum_partials_floats.argtypes = [C.c_int]
    lib.hstd_colsum_partials_floats.restype = C.c_int64
    lib.hstd_last_error.restype = C.c_char_p
    return lib


def _both(lib, x):
    """(no-alloc result, workspace result) for the bf16 matrix x, on the current stream."""
    M, N = x.shape
    stream = torch.cuda.current_stream().cuda_stream
    got = torch.full((N,), float("nan"), dtype=torch.bfloat16, device="cuda")
    want = torch.empty(N, dtype=torch.bfloat16, device="cuda")
    ws = torch.empty(lib.hstd_colsum_partials_floats(N), dtype=torch.float32, device="cuda")
    before = torch.cuda.memory_allocated()
    rc = lib.hstd_colsum_noalloc(x.data_ptr(), x.stride(0), M, N, got.data_ptr(), stream)
    assert rc == 0, lib.hstd_last_error()
    assert torch.cuda.memory_allocated() == before  # nothing was taken from torch's pool either
    assert lib.hstd_colsum(x.data_ptr(), x.stride(0), M, N, want.data_ptr(), ws.data_ptr(), stream) == 0
    torch.cuda.synchronize()
    return got, want


def test_matches_workspace_entry_point_on_two_calls_in_a_row():
    lib = _lib()
    torch.manual_seed(0)
    wide = torch.randn(4096, 2304, device="cuda").bfloat16()
    narrow = torch.randn(1000, 768, device="cuda").bfloat16() + 0.5
    for x in (wide, narrow):  # second call reuses scratch that still holds the wide call's partials
        got, want = _both(lib, x)
        ref = x.float().sum(0)
        # same kernels; the two entry points split the rows into different numbers of groups, so
        # the fp32 partial sums may differ in the last bit before rounding to bf16
        assert ((got.float() - want.float()).abs() <= 0.5 + 1e-2 * want.float().abs()).all()
        err = (got.float() - ref).abs()
        assert (err <= 0.5 + 1e-2 * ref.abs()).all(), err.max().item()  # tolerance of test_colsum


def test_rejects_widths_beyond_the_builtin_scratch():
    lib = _lib()
    max_n = lib.hstd_colsum_noalloc_max_n()
    assert max_n == 4096
    x = torch.zeros(8, max_n + 8, dtype=torch.bfloat16, device="cuda")
    out = torch.zeros(max_n + 8, dtype=torch.bfloat16, device="cuda")
    rc = lib.hstd_colsum_noalloc(x.data_ptr(), x.stride(0), 8, max_n + 8, out.data_ptr(),
                                 torch.cuda.current_stream().cuda_stream)
    assert rc != 0 and b"exceeds the built-in scratch" in lib.hstd_last_error()
    # the limit itself is served
    x = torch.ones(64, max_n, dtype=torch.bfloat16, device="cuda")
    got, want = _both(lib, x)
    assert (got.float() == 64).all() and (want.float() == 64).all()


def test_refuses_misaligned_input_instead_of_faulting():
    """Rows are read eight bf16 at a time; a caller working with raw pointers gets an error code for
    an odd base address or row stride, not a misaligned-address fault that would poison the context."""
    lib = _lib()
    x = torch.ones(16, 72, dtype=torch.bfloat16, device="cuda")
    out = torch.zeros(64, dtype=torch.bfloat16, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    assert lib.hstd_colsum_noalloc(x.data_ptr() + 2, 72, 16, 64, out.data_ptr(), s) != 0
    assert b"16-byte aligned" in lib.hstd_last_error()
    assert lib.hstd_colsum_noalloc(x.data_ptr(), 68, 16, 64, out.data_ptr(), s) != 0
    assert b"multiple of 8" in lib.hstd_last_error()
    assert lib.hstd_colsum_noalloc(x.data_ptr(), 72, 16, 64, out.data_ptr(), s) == 0
    torch.cuda.synchronize()
    assert (out.float() == 16).all()
