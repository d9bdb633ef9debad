"""Single-command fine-tuning over all visible GPUs (or the CPU): capability-equivalent of
the reference's ``scripts/singe_node_train.py`` (tf.distribute.MirroredStrategy).  The file
name keeps the reference's spelling so ``entry_point="singe_node_train.py"`` still works.
``train_batch_size`` is the *global* batch here (split across replicas), there is no LR
scaling and no ``train_runtime`` line in ``train_results.txt``.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from huggingface_sagemaker_tensorflow_distributed_b200 import runner  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200 import trainer as T  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.config import parse_args  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.parallel import hvd  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.parallel.mirrored import MirroredStrategy  # noqa: E402

if __name__ == "__main__":
    cfg = parse_args(None, distributed=False)
    strategy = MirroredStrategy()
    if strategy.ensure_replicas():  # N > 1 GPUs: this process only supervised the replicas
        sys.exit(0)
    logger = T.setup_logging()
    logger.info(cfg.to_json())

    with strategy.scope():
        replicas, rid = strategy.num_replicas_in_sync, strategy.replica_id
        logger.info("MirroredStrategy: replica %d of %d", rid, replicas)
        model, train_ds, test_ds = runner.build_model_and_data(
            cfg, rank=rid, world=replicas, per_replica_train_bs=strategy.per_replica_batch(cfg.train_batch_size),
            device=None if replicas > 1 else strategy.devices[0])
        optimizer = runner.make_optimizer(cfg, cfg.learning_rate)  # no LR scaling under MirroredStrategy
        if replicas > 1:
            optimizer = hvd.DistributedOptimizer(optimizer)  # in-graph all-reduce of the reference == our fused kernel
        model.compile(optimizer=optimizer, loss=T.SparseCategoricalCrossentropy(from_logits=True),
                      metrics=[T.SparseCategoricalAccuracy()])
        callbacks = [hvd.callbacks.BroadcastGlobalVariablesCallback(0)] if replicas > 1 else None  # mirrored creation
        runner.train_eval_save(cfg, model, train_ds, test_ds, logger, callbacks=callbacks, verbose=1 if rid == 0 else 0,
                               record_runtime=False, write_files=rid == 0)
