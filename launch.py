"""Job submitter: the local 8×B200 counterpart of the reference's SageMaker ``launch.py``.

Same knobs — hyperparameters dict, ``distribution`` selector, instance type/count, entry
point — but the "cloud" is this node: ``LocalEstimator.fit()`` sets the SageMaker
environment contract and starts one process per GPU.

    python launch.py                               # defaults below
    python launch.py --distribution mpi --instance_type local_gpu8 --model distilbert-base-uncased
"""
import argparse
import json

from huggingface_sagemaker_tensorflow_distributed_b200.launcher import HuggingFace

ap = argparse.ArgumentParser()
ap.add_argument("--entry_point", default="train.py", choices=["train.py", "singe_node_train.py"])
ap.add_argument("--distribution", default="mpi", choices=["none", "mpi", "smdistributed"])
ap.add_argument("--instance_type", default="local_gpu")
ap.add_argument("--model", default="bert-large-uncased-whole-word-masking")
ap.add_argument("--hyperparameters", default="{}", help="JSON dict merged over the defaults")
ap.add_argument("--train_channel", default=None)
ap.add_argument("--test_channel", default=None)
ap.add_argument("--output_path", default="./sm_jobs")
cli = ap.parse_args()

# hyperparameters, which are passed into the training job as --key value
hyperparameters = {"epochs": 1, "train_batch_size": 8, "eval_batch_size": 2, "model_name_or_path": cli.model}
hyperparameters.update(json.loads(cli.hyperparameters))

distribution = {
    "none": None,
    "mpi": {"mpi": {"enabled": True, "custom_mpi_options": "-verbose --NCCL_DEBUG=WARN -x RDMAV_FORK_SAFE=1"}},
    "smdistributed": {"smdistributed": {"dataparallel": {"enabled": True}}},
}[cli.distribution]

instance_count = 1
estimator = HuggingFace(
    entry_point=cli.entry_point,
    source_dir="./scripts",
    instance_type=cli.instance_type,
    instance_count=instance_count,
    distribution=distribution,
    hyperparameters=hyperparameters,
    base_job_name="hf-b200-" + str(instance_count) + "node-" + cli.instance_type.replace(".", "-").replace("_", "-"),
    output_path=cli.output_path,
    debugger_hook_config=False,
)
inputs = {k: v for k, v in (("train", cli.train_channel), ("test", cli.test_channel)) if v}

if __name__ == "__main__":
    job = estimator.fit(inputs or None)
    print(json.dumps({k: job[k] for k in ("name", "model_dir", "output_data_dir", "model_data", "output_data")}, indent=1))
