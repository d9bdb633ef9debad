"""The ctypes mirrors of the C-ABI argument structs must match include/hstd_b200.h field by field: a C program
compiled against the header prints sizeof / offsetof of every struct member and is compared with ctypes.  Catches
ABI drift (a field added on one side only) without a GPU."""
import ctypes as C
import os
import re
import shutil
import subprocess

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _mirrors():
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import _lib, kernels
    from huggingface_sagemaker_tensorflow_distributed_b200.parallel import p2p

    return {"hstd_gemm_args": _lib.GemmArgs, "hstd_step_state": kernels.StepState, "hstd_ln_args": kernels.LnArgs,
            "hstd_ln_bwd_args": kernels.LnBwdArgs, "hstd_head_args": kernels.HeadArgs,
            "hstd_allreduce_args": p2p.AllreduceArgs}


def _c_fields(header: str, struct: str):
    body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (struct, struct), header, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        # "const void* a" | "float* master" | "int32_t M, d" | "void* g_peers[8]" | several pointers after one type
        for part in decl.split(","):
            m = re.search(r"([A-Za-z_][A-Za-z_0-9]*)\s*(\[\d+\])?\s*$", part.strip())
            names.append(m.group(1))
    return names


def test_ctypes_structs_match_the_header(tmp_path):
    cc = shutil.which("g++") or shutil.which("gcc")
    if cc is None:
        pytest.skip("no C compiler")
    header = open(os.path.join(REPO, "include", "hstd_b200.h")).read()
    mirrors = _mirrors()
    lines = ['#include <cstdio>', '#include <cstddef>', '#include "hstd_b200.h"', "int main() {"]
    for struct in mirrors:
        lines.append(f'  printf("{struct} size %zu\\n", sizeof({struct}));')
        for f in _c_fields(header, struct):
            lines.append(f'  printf("{struct} {f} %zu\\n", offsetof({struct}, {f}));')
    lines.append("  return 0;\n}")
    src = tmp_path / "abi.cpp"
    src.write_text("\n".join(lines))
    exe = tmp_path / "abi"
    subprocess.run([cc, "-std=c++17", "-I", os.path.join(REPO, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout
    want = {}
    for line in out.splitlines():
        s, f, v = line.split()
        want.setdefault(s, {})[f] = int(v)
    for struct, cls in mirrors.items():
        c_names = [f for f in want[struct] if f != "size"]
        py_names = [n for n, *_ in cls._fields_]
        assert py_names == c_names, (struct, py_names, c_names)
        assert C.sizeof(cls) == want[struct]["size"], (struct, C.sizeof(cls), want[struct]["size"])
        for n in py_names:
            assert getattr(cls, n).offset == want[struct][n], (struct, n, getattr(cls, n).offset, want[struct][n])
