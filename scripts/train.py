"""Multi-process data-parallel fine-tuning: one process per B200, launched by ``launch.py``
/ ``torchrun`` / ``mpirun``.  Capability-equivalent of the reference's
``scripts/train.py`` (Horovod | smdistributed.dataparallel): same flags, same
``SM_*`` environment defaults, per-rank ``train_batch_size``, ``lr × size`` scaling,
gradient averaging across ranks, rank-0 broadcast of the initial state, ``train_results.txt``
/ ``eval_results.txt`` and a ``save_pretrained`` checkpoint.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import huggingface_sagemaker_tensorflow_distributed_b200.parallel.hvd as hvd  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200 import runner  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200 import trainer as T  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.config import parse_args  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.optim import scale_learning_rate  # noqa: E402

# process group + one-GPU-per-process pinning happen at import time, like the reference
hvd.init()


def main(argv=None):
    cfg = parse_args(argv, distributed=True)
    logger = T.setup_logging()
    logger.info("%s front-end, rank %d/%d (local %d): %s", hvd.backend_name(), hvd.rank(), hvd.size(),
                hvd.local_rank(), cfg.to_json())

    model, train_ds, test_ds = runner.build_model_and_data(
        cfg, rank=hvd.rank(), world=hvd.size(), per_replica_train_bs=cfg.train_batch_size)

    optimizer = hvd.DistributedOptimizer(
        runner.make_optimizer(cfg, scale_learning_rate(cfg.learning_rate, hvd.size())))
    model.compile(optimizer=optimizer, loss=T.SparseCategoricalCrossentropy(from_logits=True),
                  metrics=[T.SparseCategoricalAccuracy()], experimental_run_tf_function=False)

    callbacks = [hvd.callbacks.BroadcastGlobalVariablesCallback(0)]
    # only rank 0 writes result files (the reference lets every rank overwrite them)
    runner.train_eval_save(cfg, model, train_ds, test_ds, logger, callbacks=callbacks,
                           verbose=1 if hvd.rank() == 0 else 0, record_runtime=True, write_files=hvd.rank() == 0)
    hvd.shutdown()


if __name__ == "__main__":
    main()
