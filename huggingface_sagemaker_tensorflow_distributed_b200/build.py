"""Builds libhstd_b200.so (all sm_100a kernels + the C++ runtime pieces) in-tree with nvcc.

The library is plain C ABI (see include/hstd_b200.h); Python binds it with ctypes, so no
torch headers are involved and a full rebuild takes well under a minute.  Objects go to
``build/`` (git-ignored); the shared object lands in ``<package>/lib/`` so it travels with a
repo snapshot to the GPU box.

    python -m huggingface_sagemaker_tensorflow_distributed_b200.build [--force] [--verbose]
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
REPO = PKG_DIR.parent
CSRC = REPO / "csrc"
INCLUDE = REPO / "include"
BUILD_DIR = REPO / "build"
LIB_PATH = PKG_DIR / "lib" / "libhstd_b200.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC",
    "-Xcompiler", "-fvisibility=hidden",
    "-DHSTD_BUILDING_LIB",
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def _sources() -> list[Path]:
    # *.cu: sm_100a kernels; *.cpp: the host runtime (tokenizer, bucket manager, batch loader)
    return sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cpp")))


def _digest(src: Path) -> str:
    h = hashlib.sha1()
    h.update(" ".join(NVCC_FLAGS).encode())
    h.update(src.read_bytes())
    for hdr in sorted(list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))):
        h.update(hdr.read_bytes())
    return h.hexdigest()


def _compile_one(nvcc: str, src: Path, verbose: bool) -> tuple[Path, str]:
    obj = BUILD_DIR / (src.stem + ".o")
    stamp = BUILD_DIR / (src.stem + ".sha1")
    dig = _digest(src)
    if obj.exists() and stamp.exists() and stamp.read_text() == dig:
        return obj, ""
    cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{proc.stdout}\n{proc.stderr}")
    stamp.write_text(dig)
    return obj, proc.stderr


def build(force: bool = False, verbose: bool = False, watchdog: bool = False) -> Path:
    """Compile every csrc/*.cu for sm_100a and link libhstd_b200.so.  Returns its path.
    ``watchdog=True`` (development) compiles the mbarrier waits with a 2-second trap, so a barrier
    protocol bug fails a test instead of hanging the GPU."""
    nvcc = find_nvcc()
    if watchdog and "-DHSTD_MBAR_WATCHDOG" not in NVCC_FLAGS:
        NVCC_FLAGS.append("-DHSTD_MBAR_WATCHDOG")
    BUILD_DIR.mkdir(exist_ok=True)
    LIB_PATH.parent.mkdir(exist_ok=True)
    if force:
        for f in BUILD_DIR.glob("*.sha1"):
            f.unlink()
    srcs = _sources()
    if not srcs:
        raise RuntimeError(f"no CUDA sources under {CSRC}")
    objs: list[Path] = []
    logs: list[str] = []
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        for obj, log in ex.map(lambda s: _compile_one(nvcc, s, verbose), srcs):
            objs.append(obj)
            logs.append(log)
    newest = max(o.stat().st_mtime for o in objs)
    link_stamp = BUILD_DIR / "link.stamp"  # the set of objects + flags the current .so was linked from
    link_sig = " ".join(sorted(o.name for o in objs)) + " | " + " ".join(NVCC_FLAGS)
    relink = not link_stamp.exists() or link_stamp.read_text() != link_sig
    if force or relink or not LIB_PATH.exists() or LIB_PATH.stat().st_mtime < newest:
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB_PATH), *map(str, objs),
               "-lpthread", "-ldl"]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError(f"link failed:\n{proc.stdout}\n{proc.stderr}")
        link_stamp.write_text(link_sig)
    if verbose:
        sys.stderr.write("\n".join(l for l in logs if l))
    return LIB_PATH


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--watchdog", action="store_true", help="development: trap when an mbarrier wait exceeds ~2 s")
    a = ap.parse_args()
    print(build(force=a.force, verbose=a.verbose, watchdog=a.watchdog))


if __name__ == "__main__":
    main()
