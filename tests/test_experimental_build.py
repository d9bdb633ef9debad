"""The kernels under csrc/experimental/ are not part of the library, but they must keep compiling for
sm_100a without register spills worth the name (they share headers with the production kernels)."""
import os
import re
import shutil
import subprocess
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_experimental_kernels_compile_for_sm100a():
    if not (shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc")):
        pytest.skip("nvcc not available")
    r = subprocess.run([sys.executable, os.path.join(REPO, "tools", "build_experimental.py")], capture_output=True, text=True,
                       timeout=900, cwd=REPO)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "attention_fwd_pp.cu: rc=0" in r.stdout and "attention_bwd_w16.cu: rc=0" in r.stdout
    spills = [int(x) for x in re.findall(r"(\d+) bytes spill stores", r.stdout)]
    regs = [int(x) for x in re.findall(r"Used (\d+) registers", r.stdout)]
    assert spills and max(spills) <= 16, spills
    assert max(regs) <= 120, regs  # 544 threads per CTA leave 120 registers per thread
