import numpy as np

from huggingface_sagemaker_tensorflow_distributed_b200.data import (IGNORE_LABEL, BatchedDataset, build_datasets,
                                                                      load_channel, synthetic_imdb)
from huggingface_sagemaker_tensorflow_distributed_b200.models.configuration import resolve_config


def test_synthetic_is_imdb_shaped_and_deterministic():
    cfg = resolve_config("distilbert-base-uncased")
    a = synthetic_imdb(cfg, 200, 512, seed=1)
    b = synthetic_imdb(cfg, 200, 512, seed=1)
    assert a.input_ids.shape == (200, 512) and a.attention_mask.shape == (200, 512) and a.labels.shape == (200,)
    assert np.array_equal(a.input_ids, b.input_ids) and set(np.unique(a.labels)) <= {0, 1}
    # padded to max length with pad id 0 and mask 0 (scripts/train.py:80-83)
    assert ((a.attention_mask == 0) <= (a.input_ids == cfg.pad_token_id)).all()
    assert a.input_ids.max() < cfg.vocab_size and (a.input_ids[:, 0] == 101).all()
    test = synthetic_imdb(cfg, 200, 512, seed=1, split="test")
    assert not np.array_equal(a.input_ids, test.input_ids)


def test_batching_shard_and_reference_no_shard():
    cfg = resolve_config("tiny-bert")
    ds = synthetic_imdb(cfg, 50, 32)
    full = BatchedDataset(ds, 8)  # like tf.data .batch(8): 7 batches, last one short
    assert len(full) == 7
    batches = list(full)
    feats, labels = batches[-1]
    assert feats["input_ids"].shape == (8, 32) and (labels[2:] == IGNORE_LABEL).all() and (labels[:2] >= 0).all()
    r0, r1 = BatchedDataset(ds, 8, 0, 2), BatchedDataset(ds, 8, 1, 2)
    assert r0.num_samples + r1.num_samples == 50
    assert not np.intersect1d(r0.indices, r1.indices).size
    ns = BatchedDataset(ds, 8, 1, 2, shard=False)  # reference: every rank sees everything
    assert len(ns) == 7
    assert len(BatchedDataset(ds, 8, max_batches=3)) == 3


def test_channel_npz_roundtrip(tmp_path):
    cfg = resolve_config("tiny-bert")
    ds = synthetic_imdb(cfg, 10, 16)
    np.savez(tmp_path / "train.npz", input_ids=ds.input_ids, attention_mask=ds.attention_mask, label=ds.labels)
    got = load_channel(str(tmp_path), cfg, 32)
    assert got.input_ids.shape == (10, 32) and (got.input_ids[:, 16:] == cfg.pad_token_id).all()
    train, test = build_datasets(cfg, train_samples=5, test_samples=6, seq_len=32, seed=0, train_channel=str(tmp_path))
    assert len(train) == 10 and len(test) == 6


def test_uneven_shards_take_the_same_number_of_steps():
    """65 samples over 8 ranks at batch 8: rank 0 holds 9 samples, the others 8.  Every rank must still
    iterate the same number of batches (the gradient all-reduce is a collective); the short ranks pad
    with ignored rows.  Both the Python iterator and the native prefetching loader."""
    cfg = resolve_config("tiny-bert")
    ds = synthetic_imdb(cfg, 65, 16)
    for prefetch in (False, True):
        lens, seen = [], 0
        for r in range(8):
            b = BatchedDataset(ds, 8, r, 8, prefetch=prefetch)
            batches = [(f["input_ids"].clone(), l.clone()) for f, l in b]
            assert len(batches) == len(b)
            lens.append(len(batches))
            seen += sum(int((l != IGNORE_LABEL).sum()) for _, l in batches)
            if r > 0:  # the extra batch of a short rank is all padding
                assert (batches[-1][1] == IGNORE_LABEL).all()
        assert lens == [2] * 8 and seen == 65


def test_hf_datasets_route_like_the_reference(tmp_path):
    """load_dataset -> tokenizer(truncation, padding) -> dense [N, max_len] tables (scripts/train.py:69-100), from a
    local DatasetDict (no network here): train split padded to the longest example, test to max_length, both
    densified to the model's sequence length."""
    import datasets as hfds

    from huggingface_sagemaker_tensorflow_distributed_b200.data import try_load_tokenizer

    dd = hfds.DatasetDict({
        "train": hfds.Dataset.from_dict({"text": ["a fine film", "dull and far too long , really", "ok"], "label": [1, 0, 1]}),
        "test": hfds.Dataset.from_dict({"text": ["not bad at all", "awful"], "label": [1, 0]}),
    })
    dd.save_to_disk(str(tmp_path / "imdb_local"))
    cfg = resolve_config("tiny-bert")
    tok = try_load_tokenizer("tiny-bert")
    if tok is None:
        import pytest
        pytest.skip("native tokenizer library not built")
    train, test = build_datasets(cfg, train_samples=-1, test_samples=-1, seq_len=32, seed=0, tokenizer=tok,
                                 data_source="hf:" + str(tmp_path / "imdb_local"))
    assert train.input_ids.shape == (3, 32) and test.input_ids.shape == (2, 32)
    assert list(train.labels) == [1, 0, 1] and list(test.labels) == [1, 0]
    assert (train.attention_mask.sum(1) >= 3).all() and train.input_ids.max() < cfg.vocab_size
    assert ((train.attention_mask == 0) <= (train.input_ids == cfg.pad_token_id)).all()
    assert train.source.startswith("hf:")
