"""bench.py contract pieces that can be checked without a GPU."""
import json
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_reports_unavailable_and_exits_zero():
    """The reference (TensorFlow 2.4 + Horovod / smdistributed scripts) cannot be installed on this
    image; the reference arm must say so in ONE JSON line and exit 0."""
    r = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "2",
                        "--warmup", "1"], capture_output=True, text=True, timeout=300, cwd=REPO)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference"
    if "unavailable" in d:
        assert isinstance(d["unavailable"], str) and "\n" not in d["unavailable"] and d["unavailable"]
    else:  # a box where the reference does run: the full line is owed
        for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step"):
            assert k in d


def test_bench_cli_defaults_are_the_baseline_config():
    sys.path.insert(0, REPO)
    import importlib

    bench = importlib.import_module("bench")
    a = bench.parse([])
    assert a.gpus == 1 and a.steps > 0 and a.warmup >= 3
    assert a.model == "distilbert-base-uncased" and a.seq == 512 and a.batch == 8
