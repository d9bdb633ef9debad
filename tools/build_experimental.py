"""Compile-check of the kernels under csrc/experimental/ for sm_100a (objects only, into build/experimental/;
they are NOT linked into libhstd_b200.so).  Prints registers / stack / spills per kernel."""
import subprocess
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
from huggingface_sagemaker_tensorflow_distributed_b200.build import CSRC, INCLUDE, NVCC_FLAGS, find_nvcc  # noqa: E402

out = REPO / "build" / "experimental"
out.mkdir(parents=True, exist_ok=True)
rc = 0
for src in sorted((CSRC / "experimental").glob("*.cu")):
    obj = out / (src.stem + ".o")
    cmd = [find_nvcc(), *NVCC_FLAGS, "-Xptxas", "-v", "-I", str(INCLUDE), "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
    p = subprocess.run(cmd, capture_output=True, text=True)
    print(f"== {src.name}: rc={p.returncode}")
    if p.returncode != 0:
        print(p.stdout[-3000:], p.stderr[-6000:])
        rc = 1
        continue
    lines = [l for l in p.stderr.splitlines() if "Used" in l or "spill" in l or "Compiling entry" in l]
    for i in range(0, len(lines)):
        print("  ", lines[i].replace("ptxas info    : ", "")[:160])
sys.exit(rc)
