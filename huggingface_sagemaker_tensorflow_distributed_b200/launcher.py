"""Local estimator-style job launcher.

The reference builds a SageMaker ``HuggingFace`` estimator and calls ``.fit()``
(launch.py:36-55), which provisions instances, sets the ``SM_*`` environment, turns the
hyperparameter dict into argv and starts the entry script under ``python``, ``mpirun``
(Horovod) or the smddp runner depending on ``distribution`` (launch.py:19-24).  There is no
cloud here: ``LocalEstimator`` keeps the same constructor shape and does the container's job
on one 8×B200 node — one process per GPU with ``RANK / LOCAL_RANK / WORLD_SIZE /
MASTER_*`` set, stdout prefixed per rank, the whole job killed (non-zero exit) if any rank
dies, and ``model.tar.gz`` / ``output.tar.gz`` produced at the end like SageMaker's upload.
"""
from __future__ import annotations

import os
import signal
import socket
import subprocess
import sys
import tarfile
import threading
import time
from typing import Any, Dict, List, Optional

from .config import hyperparameters_to_argv, sagemaker_env

# instance types the reference names (launch.py:26-28) → GPUs per instance; "local_gpu*"
# are this framework's own names for a B200 node.
INSTANCE_GPUS = {
    "ml.p3.2xlarge": 1, "ml.p3.8xlarge": 4, "ml.p3.16xlarge": 8, "ml.p3dn.24xlarge": 8, "ml.p4d.24xlarge": 8,
    "local": 0, "local_cpu": 0, "local_gpu": -1, "local_gpu1": 1, "local_gpu2": 2, "local_gpu4": 4, "local_gpu8": 8,
}


def _visible_gpus() -> int:
    try:
        import torch

        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:  # noqa: BLE001
        return 0


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def distribution_mode(distribution: Optional[dict]) -> str:
    """'none' | 'mpi' (Horovod launch) | 'smddp' — the three selectors of launch.py:19-24."""
    if not distribution:
        return "none"
    if distribution.get("smdistributed", {}).get("dataparallel", {}).get("enabled"):
        return "smddp"
    if distribution.get("mpi", {}).get("enabled"):
        return "mpi"
    return "none"


def parse_custom_mpi_options(opts: str) -> Dict[str, str]:
    """``-x NAME=VALUE`` and ``--NCCL_DEBUG=INFO``-style options of
    ``custom_mpi_options`` (launch.py:22) become environment variables of the ranks."""
    env: Dict[str, str] = {}
    toks = (opts or "").split()
    i = 0
    while i < len(toks):
        t = toks[i]
        if t == "-x" and i + 1 < len(toks):
            if "=" in toks[i + 1]:
                k, v = toks[i + 1].split("=", 1)
                env[k] = v
            i += 2
            continue
        if t.startswith("--") and "=" in t:
            k, v = t[2:].split("=", 1)
            env[k] = v
        i += 1
    return env


def spawn_ranks(argv: List[str], base_env: Dict[str, str], nproc: int, port: int):
    """Start ``nproc`` copies of ``argv`` with torchrun-style rank environment; stdout/stderr of
    every rank is forwarded line by line, prefixed with ``[rank]`` when nproc > 1."""
    procs: List[subprocess.Popen] = []
    threads: List[threading.Thread] = []

    def pump(rank: int, pipe) -> None:
        for line in iter(pipe.readline, ""):
            sys.stdout.write(f"[{rank}] {line}" if nproc > 1 else line)
            sys.stdout.flush()
        pipe.close()

    for r in range(nproc):
        env = dict(base_env)
        env.update({"RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(nproc), "LOCAL_WORLD_SIZE": str(nproc),
                    "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)})
        p = subprocess.Popen(argv, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                             start_new_session=True)
        procs.append(p)
        t = threading.Thread(target=pump, args=(r, p.stdout), daemon=True)
        t.start()
        threads.append(t)
    return procs, threads


def wait_all_or_kill(procs: List[subprocess.Popen], timeout_s: Optional[float] = None) -> int:
    """Failure detection: the first non-zero exit (or the timeout) kills every other rank.
    Returns the job's exit code (0 only if every rank exited 0)."""
    t0 = time.time()
    alive = set(range(len(procs)))
    rc_final = 0
    while alive:
        for i in list(alive):
            rc = procs[i].poll()
            if rc is None:
                continue
            alive.discard(i)
            if rc != 0 and rc_final == 0:
                rc_final = rc
                for j in alive:
                    _kill_group(procs[j])
        if timeout_s is not None and time.time() - t0 > timeout_s and alive:
            rc_final = rc_final or 124
            for j in alive:
                _kill_group(procs[j])
        time.sleep(0.05)
    return rc_final


class LocalEstimator:
    def __init__(self, entry_point: str, source_dir: str = "./scripts", instance_type: str = "local_gpu",
                 instance_count: int = 1, hyperparameters: Optional[Dict[str, Any]] = None,
                 distribution: Optional[dict] = None, base_job_name: Optional[str] = None,
                 output_path: str = "./sm_jobs", role: Optional[str] = None, session: Any = None,
                 image_uri: Optional[str] = None, py_version: Optional[str] = None,
                 transformers_version: Optional[str] = None, tensorflow_version: Optional[str] = None,
                 debugger_hook_config: Any = None, volume_size: Optional[int] = None,
                 environment: Optional[Dict[str, str]] = None, num_gpus: Optional[int] = None,
                 python_executable: Optional[str] = None, timeout_s: Optional[float] = None):
        if instance_count != 1:
            raise ValueError("multi-node is not exercised by the reference (launch.py:29) nor supported here")
        self.entry_point, self.source_dir = entry_point, source_dir
        self.instance_type, self.instance_count = instance_type, instance_count
        self.hyperparameters = dict(hyperparameters or {})
        self.distribution = distribution
        self.base_job_name = base_job_name or ("hf-b200-" + str(instance_count) + "node-" +
                                               instance_type.replace(".", "-").replace("_", "-"))
        self.output_path = output_path
        self.environment = dict(environment or {})
        self.num_gpus = num_gpus
        self.python = python_executable or sys.executable
        self.timeout_s = timeout_s
        # accepted for signature parity with sagemaker.huggingface.HuggingFace; unused locally
        self._unused = dict(role=role, session=session, image_uri=image_uri, py_version=py_version,
                            transformers_version=transformers_version, tensorflow_version=tensorflow_version,
                            debugger_hook_config=debugger_hook_config, volume_size=volume_size)
        self.latest_job: Optional[dict] = None

    # ------------------------------------------------------------------
    def _gpus_for_instance(self) -> int:
        if self.num_gpus is not None:
            return self.num_gpus
        n = INSTANCE_GPUS.get(self.instance_type, -1)
        vis = _visible_gpus()
        if n < 0:
            return vis
        return min(n, vis) if vis else 0

    def plan(self) -> dict:
        """What ``fit`` would run: process count, per-rank env and argv."""
        mode = distribution_mode(self.distribution)
        gpus = self._gpus_for_instance()
        nproc = max(1, gpus) if mode in ("mpi", "smddp") else 1
        script = os.path.join(self.source_dir, self.entry_point)
        argv = [self.python, script] + hyperparameters_to_argv(self.hyperparameters)
        return {"mode": mode, "gpus": gpus, "nproc": nproc, "argv": argv}

    def fit(self, inputs: Optional[Dict[str, str]] = None, wait: bool = True, job_name: Optional[str] = None) -> dict:
        plan = self.plan()
        job_name = job_name or f"{self.base_job_name}-{time.strftime('%Y-%m-%d-%H-%M-%S')}"
        job_dir = os.path.abspath(os.path.join(self.output_path, job_name))
        model_dir, out_dir = os.path.join(job_dir, "model"), os.path.join(job_dir, "output", "data")
        os.makedirs(model_dir, exist_ok=True)
        os.makedirs(out_dir, exist_ok=True)
        channels = {k: os.path.abspath(v) for k, v in (inputs or {}).items()}
        base_env = dict(os.environ)
        base_env.update(sagemaker_env(model_dir=model_dir, output_data_dir=out_dir, num_gpus=plan["gpus"],
                                      hyperparameters=self.hyperparameters, channels=channels,
                                      smddp=plan["mode"] == "smddp", job_name=job_name))
        if plan["mode"] == "mpi":
            base_env.update(parse_custom_mpi_options(self.distribution["mpi"].get("custom_mpi_options", "")))
        base_env.update(self.environment)
        nproc = plan["nproc"]
        procs, threads = spawn_ranks(plan["argv"], base_env, nproc, _free_port())
        self.latest_job = {"name": job_name, "dir": job_dir, "model_dir": model_dir, "output_data_dir": out_dir,
                           "procs": procs, "plan": plan}
        if not wait:
            return self.latest_job
        rc = self._wait(procs)
        for t in threads:
            t.join(timeout=5)
        self.latest_job["returncode"] = rc
        if rc != 0:
            raise RuntimeError(f"training job {job_name} failed: a rank exited with code {rc}")
        # SageMaker tars SM_MODEL_DIR / SM_OUTPUT_DATA_DIR to S3 at job end
        self.latest_job["model_data"] = _tar(model_dir, os.path.join(job_dir, "model.tar.gz"))
        self.latest_job["output_data"] = _tar(out_dir, os.path.join(job_dir, "output.tar.gz"))
        return self.latest_job

    def _wait(self, procs: List[subprocess.Popen]) -> int:
        return wait_all_or_kill(procs, self.timeout_s)


def _kill_group(p: subprocess.Popen) -> None:
    try:
        os.killpg(os.getpgid(p.pid), signal.SIGTERM)
    except (ProcessLookupError, PermissionError):
        pass


def _tar(src_dir: str, dst: str) -> str:
    with tarfile.open(dst, "w:gz") as tf:
        for name in sorted(os.listdir(src_dir)):
            tf.add(os.path.join(src_dir, name), arcname=name)
    return dst


# name parity with ``from sagemaker.huggingface import HuggingFace`` (launch.py:4)
HuggingFace = LocalEstimator
