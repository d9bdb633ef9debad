"""Single-process fine-tuning over all visible GPUs (or the CPU): capability-equivalent of
the reference's ``scripts/singe_node_train.py`` (tf.distribute.MirroredStrategy).  The file
name keeps the reference's spelling so ``entry_point="singe_node_train.py"`` still works.
``train_batch_size`` is the *global* batch here, there is no LR scaling and no
``train_runtime`` line in ``train_results.txt``.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from huggingface_sagemaker_tensorflow_distributed_b200 import runner  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200 import trainer as T  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.config import parse_args  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.parallel.mirrored import MirroredStrategy  # noqa: E402

if __name__ == "__main__":
    cfg = parse_args(None, distributed=False)
    logger = T.setup_logging()
    logger.info(cfg.to_json())

    strategy = MirroredStrategy()
    with strategy.scope():
        replicas = strategy.num_replicas_in_sync
        logger.info("MirroredStrategy with %d replica(s): %s", replicas, [str(d) for d in strategy.devices])
        model, train_ds, test_ds = runner.build_model_and_data(
            cfg, rank=0, world=1, per_replica_train_bs=cfg.train_batch_size, device=strategy.devices[0])
        model.compile(optimizer=runner.make_optimizer(cfg, cfg.learning_rate),
                      loss=T.SparseCategoricalCrossentropy(from_logits=True), metrics=[T.SparseCategoricalAccuracy()])
        runner.train_eval_save(cfg, model, train_ds, test_ds, logger, verbose=1, record_runtime=False)
