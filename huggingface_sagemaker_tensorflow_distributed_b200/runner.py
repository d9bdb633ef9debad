"""Shared body of the two entry scripts.

``scripts/train.py`` (multi-process data parallel, Horovod/smddp-shaped) and
``scripts/singe_node_train.py`` (single process, MirroredStrategy-shaped) differ only in
how they distribute; everything else — tokenizer/dataset build, compile, fit, results
files, evaluate, save — is this module.
"""
from __future__ import annotations

import os
import time

from . import trainer as T
from .config import TrainConfig
from .data import BatchedDataset, build_datasets
from .models.configuration import resolve_config
from .optim import Adam
from .parallel import comm


def build_model_and_data(cfg: TrainConfig, *, rank: int, world: int, per_replica_train_bs: int, device=None):
    mcfg = resolve_config(cfg.model_name_or_path)
    seq_len = cfg.max_seq_length or mcfg.model_max_length
    seq_len = min(seq_len, mcfg.model_max_length)
    model = T.AutoModelForSequenceClassification.from_pretrained(
        cfg.model_name_or_path, device=device, dtype=cfg.dtype, seed=cfg.seed, dropout=cfg.dropout,
        **({} if (device is not None and device.type == "cpu") or comm.device().type == "cpu" else
           {"bucket_mb": cfg.bucket_mb, "comm_mode": cfg.comm, "use_cuda_graph": cfg.cuda_graph}))
    train, test = build_datasets(model.cfg, train_samples=cfg.train_samples, test_samples=cfg.test_samples,
                                 seq_len=seq_len, seed=cfg.seed, train_channel=cfg.train_channel,
                                 test_channel=cfg.test_channel, tokenizer=model.tokenizer,
                                 data_source=cfg.data_source)
    pin = model.backend.device.type == "cuda"
    train_ds = BatchedDataset(train, per_replica_train_bs, rank, world, shard=not cfg.no_shard,
                              max_batches=cfg.max_steps, pin_memory=pin, prefetch=pin and cfg.prefetch)
    # the reference evaluates the full test set on every rank (scripts/train.py:170)
    test_ds = BatchedDataset(test, cfg.eval_batch_size, 0, 1, shard=False, max_batches=cfg.max_eval_steps,
                             pin_memory=pin, prefetch=pin and cfg.prefetch)
    return model, train_ds, test_ds


def make_optimizer(cfg: TrainConfig, learning_rate: float) -> Adam:
    return Adam(learning_rate=learning_rate, epsilon=cfg.adam_epsilon, weight_decay=cfg.weight_decay,
                lr_schedule=cfg.lr_schedule, warmup_steps=cfg.warmup_steps)


def train_eval_save(cfg: TrainConfig, model, train_ds, test_ds, logger, *, callbacks=None, verbose: int = 1,
                    record_runtime: bool = True, write_files: bool = True) -> dict:
    out = {}
    model.set_metrics_sink(cfg.metrics_jsonl)
    if cfg.do_train:
        logger.info("*** Train ***")
        t0 = time.time()
        hist = model.fit(train_ds, epochs=cfg.epochs, callbacks=callbacks, verbose=verbose,
                         batch_size=cfg.train_batch_size, validation_batch_size=cfg.eval_batch_size,
                         log_every=cfg.log_every)
        runtime = time.time() - t0
        out["history"], out["train_runtime"] = hist.history, runtime
        logger.info("train_runtime = %.4f s, %.1f samples/s (global)", runtime,
                    model.last_fit_stats.get("train_samples_per_second", 0.0))
        if write_files:
            logger.info("***** Train results *****")
            for k, v in hist.history.items():
                logger.info("  %s = %s", k, v)
            T.write_train_results(os.path.join(cfg.output_data_dir, "train_results.txt"), hist,
                                  runtime if record_runtime else None)
    if cfg.do_eval:
        logger.info("*** Evaluate ***")
        result = model.evaluate(test_ds, batch_size=cfg.eval_batch_size, return_dict=True)
        out["eval"] = result
        if write_files:
            logger.info("***** Eval results *****")
            for k, v in result.items():
                logger.info("  %s = %s", k, v)
            T.write_eval_results(os.path.join(cfg.output_data_dir, "eval_results.txt"), result)
    # unconditional, as in the reference (scripts/train.py:182-183) — but rank 0 + barrier
    model.save_pretrained(cfg.model_dir, save_training_state=cfg.save_training_state)
    return out
