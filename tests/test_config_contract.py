"""Argv / environment contract (reference scripts/train.py:36-52, launch.py:13-24)."""
import json

import pytest

from huggingface_sagemaker_tensorflow_distributed_b200 import config as C


def test_reference_flag_defaults(monkeypatch):
    for k in ("SM_OUTPUT_DATA_DIR", "SM_MODEL_DIR", "SM_NUM_GPUS"):
        monkeypatch.delenv(k, raising=False)
    cfg = C.parse_args(["--model_name_or_path", "distilbert-base-uncased"])
    assert (cfg.epochs, cfg.train_batch_size, cfg.eval_batch_size) == (3, 8, 4)
    assert cfg.learning_rate == pytest.approx(5e-5) and isinstance(cfg.learning_rate, float)
    assert cfg.do_train and cfg.do_eval
    # the reference raises KeyError when SM_* is unset; we fall back
    assert cfg.output_data_dir == "./output" and cfg.model_dir == "./model" and cfg.n_gpus is None
    assert cfg.weight_decay == 0.0 and cfg.lr_schedule == "constant"  # == the reference's plain Adam


def test_env_defaults_and_unknown_flags(monkeypatch):
    monkeypatch.setenv("SM_OUTPUT_DATA_DIR", "/opt/ml/output/data")
    monkeypatch.setenv("SM_MODEL_DIR", "/opt/ml/model")
    monkeypatch.setenv("SM_NUM_GPUS", "8")
    monkeypatch.setenv("SM_CHANNEL_TRAIN", "/opt/ml/input/data/train")
    cfg = C.parse_args(["--model_name_or_path", "x", "--some_unknown_flag", "1", "--learning_rate", "3e-5"])
    assert cfg.output_data_dir == "/opt/ml/output/data" and cfg.model_dir == "/opt/ml/model" and cfg.n_gpus == "8"
    assert cfg.train_channel == "/opt/ml/input/data/train"
    assert cfg.learning_rate == pytest.approx(3e-5)
    assert "--some_unknown_flag" in cfg.extra["unknown_args"]


def test_single_node_parser_has_no_n_gpus():
    p = C.build_parser(distributed=False)
    assert "--n_gpus" not in p._option_string_actions
    assert "--n_gpus" in C.build_parser(distributed=True)._option_string_actions


@pytest.mark.parametrize("s,v", [("False", False), ("false", False), ("0", False), ("True", True), ("yes", True)])
def test_bool_flags_parse_properly(s, v):
    assert C.parse_args(["--do_train", s, "--model_name_or_path", "x"]).do_train is v


def test_hyperparameters_to_argv_matches_launch_dict():
    hp = {"epochs": 1, "train_batch_size": 8, "eval_batch_size": 2,
          "model_name_or_path": "bert-large-uncased-whole-word-masking"}
    argv = C.hyperparameters_to_argv(hp)
    assert argv == ["--epochs", "1", "--train_batch_size", "8", "--eval_batch_size", "2", "--model_name_or_path",
                    "bert-large-uncased-whole-word-masking"]
    cfg = C.parse_args(argv)
    assert cfg.epochs == 1 and cfg.eval_batch_size == 2 and cfg.learning_rate == pytest.approx(5e-5)


def test_sagemaker_env_and_smddp_switch():
    env = C.sagemaker_env(model_dir="/m", output_data_dir="/o/data", num_gpus=8, hyperparameters={"epochs": 1},
                          channels={"train": "/d/train"}, smddp=True)
    assert env["SM_MODEL_DIR"] == "/m" and env["SM_NUM_GPUS"] == "8" and env["SM_CHANNEL_TRAIN"] == "/d/train"
    assert json.loads(env["SM_HPS"]) == {"epochs": 1}
    assert C.is_sagemaker_dp_enabled(env)
    assert not C.is_sagemaker_dp_enabled(C.sagemaker_env(model_dir="/m", output_data_dir="/o", num_gpus=1,
                                                          hyperparameters={}))
    assert not C.is_sagemaker_dp_enabled({"SM_FRAMEWORK_PARAMS": "not json"})
