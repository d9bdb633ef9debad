"""Estimator-shaped local launcher (reference launch.py:13-55)."""
import os
import sys
import textwrap

import pytest

from huggingface_sagemaker_tensorflow_distributed_b200.launcher import (HuggingFace, distribution_mode,
                                                                          parse_custom_mpi_options)


def test_distribution_selector_three_modes():
    assert distribution_mode(None) == "none"
    assert distribution_mode({"smdistributed": {"dataparallel": {"enabled": True}}}) == "smddp"
    assert distribution_mode({"mpi": {"enabled": True, "custom_mpi_options": "-verbose"}}) == "mpi"


def test_custom_mpi_options_become_env():
    env = parse_custom_mpi_options("-verbose --NCCL_DEBUG=INFO -x RDMAV_FORK_SAFE=1")
    assert env == {"NCCL_DEBUG": "INFO", "RDMAV_FORK_SAFE": "1"}


def test_plan_argv_and_process_count():
    hp = {"epochs": 1, "train_batch_size": 8, "eval_batch_size": 2, "model_name_or_path": "bert-large-uncased"}
    est = HuggingFace(entry_point="train.py", source_dir="./scripts", instance_type="ml.p3.16xlarge",
                      hyperparameters=hp, distribution={"mpi": {"enabled": True}}, num_gpus=8, role="r",
                      image_uri="x", py_version="py37", debugger_hook_config=False)
    plan = est.plan()
    assert plan["nproc"] == 8 and plan["mode"] == "mpi"
    assert plan["argv"][1:] == ["./scripts/train.py", "--epochs", "1", "--train_batch_size", "8", "--eval_batch_size",
                                "2", "--model_name_or_path", "bert-large-uncased"]
    assert HuggingFace(entry_point="train.py", distribution=None, num_gpus=8).plan()["nproc"] == 1
    assert est.base_job_name == "hf-b200-1node-ml-p3-16xlarge"
    with pytest.raises(ValueError):
        HuggingFace(entry_point="train.py", instance_count=2)


def _script(tmp_path, body):
    d = tmp_path / "src"
    d.mkdir(exist_ok=True)
    (d / "job.py").write_text(textwrap.dedent(body))
    return str(d)


def test_env_contract_reaches_every_rank_and_artifacts_are_tarred(tmp_path):
    src = _script(tmp_path, """
        import os, sys, json
        r = os.environ["RANK"]
        assert os.environ["WORLD_SIZE"] == "2" and os.environ["MASTER_ADDR"] == "127.0.0.1"
        assert json.loads(os.environ["SM_HPS"]) == {"epochs": 2}
        assert sys.argv[1:] == ["--epochs", "2"], sys.argv
        assert "sagemaker_distributed_dataparallel_enabled" in os.environ["SM_FRAMEWORK_PARAMS"]
        assert os.environ["SM_CHANNEL_TRAIN"].endswith("train")
        open(os.path.join(os.environ["SM_OUTPUT_DATA_DIR"], f"r{r}.txt"), "w").write("ok")
        if r == "0":
            open(os.path.join(os.environ["SM_MODEL_DIR"], "config.json"), "w").write("{}")
        print("hello from", r)
    """)
    (tmp_path / "train").mkdir()
    est = HuggingFace(entry_point="job.py", source_dir=src, hyperparameters={"epochs": 2}, num_gpus=2,
                      distribution={"smdistributed": {"dataparallel": {"enabled": True}}},
                      output_path=str(tmp_path / "jobs"))
    job = est.fit({"train": str(tmp_path / "train")})
    assert sorted(os.listdir(job["output_data_dir"])) == ["r0.txt", "r1.txt"]
    assert os.path.exists(job["model_data"]) and job["model_data"].endswith("model.tar.gz")


def test_one_dead_rank_kills_the_job(tmp_path):
    src = _script(tmp_path, """
        import os, sys, time
        if os.environ["RANK"] == "1":
            sys.exit(3)
        time.sleep(60)
    """)
    est = HuggingFace(entry_point="job.py", source_dir=src, num_gpus=2, distribution={"mpi": {"enabled": True}},
                      output_path=str(tmp_path / "jobs"), timeout_s=30)
    import time
    t0 = time.time()
    with pytest.raises(RuntimeError, match="exited with code 3"):
        est.fit()
    assert time.time() - t0 < 20


def test_mirrored_strategy_single_command_fans_out_to_replicas(tmp_path):
    """singe_node_train.py with 2 (CPU stand-in) devices: one command, two replica processes,
    global batch split in two, one set of result files (reference scripts/singe_node_train.py)."""
    import subprocess
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    wrapper = tmp_path / "run.py"
    wrapper.write_text(textwrap.dedent(f"""
        import os, sys, runpy
        sys.path.insert(0, {repo!r})
        os.environ["HSTD_FORCE_CPU"] = "1"
        import huggingface_sagemaker_tensorflow_distributed_b200.parallel.mirrored as M
        _orig = M.MirroredStrategy.__init__
        def _init(self, devices=None):
            _orig(self, devices if devices is not None else ["cpu", "cpu"])
        M.MirroredStrategy.__init__ = _init
        sys.argv = [{os.path.join(repo, 'scripts', 'singe_node_train.py')!r}] + sys.argv[1:]
        runpy.run_path(sys.argv[0], run_name="__main__")
    """))
    env = dict(os.environ, SM_OUTPUT_DATA_DIR=str(tmp_path / "out"), SM_MODEL_DIR=str(tmp_path / "model"),
               OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, str(wrapper), "--model_name_or_path", "tiny-distilbert", "--epochs", "1",
                        "--train_batch_size", "8", "--max_seq_length", "16", "--train_samples", "32", "--test_samples",
                        "8", "--log_every", "0"], env=env, capture_output=True, text=True, timeout=280)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "[0]" in r.stdout and "[1]" in r.stdout and "replica 1 of 2" in r.stdout
    assert sorted(os.listdir(tmp_path / "out")) == ["eval_results.txt", "train_results.txt"]
    assert "train_runtime" not in open(tmp_path / "out" / "train_results.txt").read()
    # model + tokenizer files, as model.save_pretrained + tokenizer.save_pretrained leave them (scripts/train.py:182-183)
    assert sorted(os.listdir(tmp_path / "model")) == ["config.json", "model.safetensors", "special_tokens_map.json",
                                                      "tokenizer_config.json", "vocab.txt"]
