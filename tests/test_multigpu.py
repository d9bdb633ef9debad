"""Collective tier: the fused P2P all-reduce kernels across 2+ GPUs —
agreement with a gathered fp32 reference, bitwise agreement across ranks, fused AdamW vs
reduce-then-update, and end-to-end data-parallel training (p2p kernels vs NCCL plumbing)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(which, n):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(HERE, "mgpu_worker.py"), which]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-3000:]
    assert "ALL OK" in r.stdout, r.stdout[-3000:]
    return r.stdout


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_collectives_all_variants():
    out = _run("collectives", min(torch.cuda.device_count(), 8) // 2 * 2 if torch.cuda.device_count() < 8 else 8)
    assert "OK allreduce two_shot" in out and "OK allreduce one_shot" in out and "OK fused_update two_shot" in out


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_data_parallel_training_p2p_matches_nccl():
    out = _run("training", 2)
    assert "OK training p2p==nccl" in out
