"""Native host runtime (CPU): WordPiece tokenizer, fusion-bucket manager, prefetching loader.

These live in libhstd_b200.so next to the kernels but touch no GPU; the library is built on
demand with nvcc (the same thing ``__graft_entry__.build()`` does).
"""
import json
import os
from types import SimpleNamespace

import numpy as np
import pytest


@pytest.fixture(scope="module", autouse=True)
def _native_lib():
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import _lib

    if not _lib.available():
        from huggingface_sagemaker_tensorflow_distributed_b200 import build

        build.build()
    return _lib.load()


def _tok(vocab_size=30522):
    from huggingface_sagemaker_tensorflow_distributed_b200.tokenization import WordPieceTokenizer

    return WordPieceTokenizer.synthetic(vocab_size)


# ------------------------------------------------------------------------------------------ tokenizer
def test_wordpiece_matches_hf_tokenizers_library(tmp_path):
    """Oracle: the `tokenizers` (Rust) BertWordPieceTokenizer the reference reaches through
    AutoTokenizer (scripts/train.py:69), on the same vocabulary."""
    tk = pytest.importorskip("tokenizers")
    from huggingface_sagemaker_tensorflow_distributed_b200.tokenization import synthetic_reviews

    tok = _tok()
    tok.save_pretrained(tmp_path)
    oracle = tk.BertWordPieceTokenizer(str(tmp_path / "vocab.txt"), lowercase=True)
    vocab = [tok.convert_ids_to_tokens(i) for i in range(1000, 9000)]
    texts, _ = synthetic_reviews(64, seed=3, vocab=vocab)
    texts += ["The movie was Greatly unbelievable, café! naïve Ünïcödé façade — “quoted” 中文 text… ok?",
              "", "   ", "tabs\tand\nnewlines\r\n  x", "a" * 150 + " tail", "semi;colon:co-op's $5.00 (really) [x] {y} <z> #tag @me",
              "Ça va? Łódź Ærøskøbing straße ĲSSEL İstanbul", "emoji 😀 and\x00nul\x07bell​zero-width"]
    ids, mask, lengths = tok._native.encode_batch(texts, 512)
    for r, text in enumerate(texts):
        want = oracle.encode(text.replace("\x00", " ")).ids
        got = ids[r, : int(mask[r].sum())].tolist()
        if len(want) > 512:  # truncation keeps [CLS] ... [SEP]
            want_row = want[:511] + [want[-1]]
        else:
            want_row = want
        assert got == want_row, (r, text[:60])
        assert lengths[r] == len(want)


def test_tokenizer_call_semantics_and_roundtrip(tmp_path):
    from huggingface_sagemaker_tensorflow_distributed_b200.tokenization import AutoTokenizer, WordPieceTokenizer

    tok = _tok(2000 + 48)
    assert (tok.pad_token_id, tok.unk_token_id, tok.cls_token_id, tok.sep_token_id) == (0, 100, 101, 102)
    assert tok.model_max_length == 512
    texts = ["a great movie", "the film was boring and the plot was a mess , really"]
    longest = tok(texts, truncation=True, padding="longest")
    assert len(longest["input_ids"][0]) == len(longest["input_ids"][1]) < 512
    assert longest["attention_mask"][0][-1] == 0 and longest["input_ids"][0][0] == 101
    full = tok(texts, truncation=True, padding="max_length", return_tensors="np")
    assert full["input_ids"].shape == (2, 512) and full["input_ids"].dtype == np.int32
    short = tok(texts, truncation=True, padding="max_length", max_length=6)
    assert [len(r) for r in short["input_ids"]] == [6, 6] and short["input_ids"][1][-1] == 102  # [SEP] survives truncation
    single = tok("a great movie")
    assert single["input_ids"][0] == 101 and single["input_ids"][-1] == 102 and all(single["attention_mask"])
    assert tok.convert_ids_to_tokens(single["input_ids"])[1:-1] == tok.tokenize("a great movie")
    assert tok.decode(single["input_ids"]) == "a great movie"
    assert tok.batch_decode(tok(["Wonderful , moving film !", "BAD"])["input_ids"]) == ["wonderful , moving film !", "bad"]
    # save_pretrained / from_pretrained (scripts/train.py:183) through the Auto class
    files = tok.save_pretrained(tmp_path)
    assert {os.path.basename(f) for f in files} == {"vocab.txt", "tokenizer_config.json", "special_tokens_map.json"}
    assert json.load(open(tmp_path / "tokenizer_config.json"))["tokenizer_class"] == "BertTokenizer"
    again = AutoTokenizer.from_pretrained(str(tmp_path))
    assert isinstance(again, WordPieceTokenizer) and again.vocab_size == tok.vocab_size
    assert again(texts, padding="longest") == longest
    # hub name, no network: synthetic vocabulary of the model's size
    auto = AutoTokenizer.from_pretrained("distilbert-base-uncased")
    assert auto.vocab_size == 30522 and "synthetic" in auto.source


def test_byte_level_bpe_matches_hf_tokenizers_library(tmp_path):
    """RoBERTa-style vocab.json + merges.txt (trained offline on synthetic reviews by the `tokenizers`
    library, which is also the oracle): same ids for ASCII, contractions, whitespace runs, Latin-1,
    CJK and emoji; <s> ... </s> framing, pad id 1, truncation keeps </s>."""
    tk = pytest.importorskip("tokenizers")
    from huggingface_sagemaker_tensorflow_distributed_b200.tokenization import (AutoTokenizer, ByteLevelBPETokenizer,
                                                                                 synthetic_reviews)

    texts, _ = synthetic_reviews(200, seed=2)
    extra = ["The movie was great, I'd say it's the best!  Really... 123 times better than café №5 — naïve?",
             "hello world\n\nnew  lines\tand tabs  ", "it's they're we've I'm you'll he'd 'quoted' 'such", "", "   ", " x", "x ",
             "a\n b", "Ünïcödé façade — “quoted” 中文 text… ok? 😀 emoji nbsp\u00a0here 3.14 1,000 x² ½"]
    oracle = tk.ByteLevelBPETokenizer()
    oracle.train_from_iterator(texts + extra * 5, vocab_size=1500, show_progress=False,
                               special_tokens=["<s>", "<pad>", "</s>", "<unk>", "<mask>"])
    oracle.save_model(str(tmp_path))
    tok = AutoTokenizer.from_pretrained(str(tmp_path))
    assert isinstance(tok, ByteLevelBPETokenizer) and tok.vocab_size == oracle.get_vocab_size()
    assert (tok.cls_token_id, tok.pad_token_id, tok.sep_token_id, tok.unk_token_id) == (0, 1, 2, 3)
    sample = texts[:60] + extra
    ids, mask, lengths = tok._native.encode_batch(sample, 2048, add_special_tokens=False)
    for r, text in enumerate(sample):
        want = oracle.encode(text).ids
        assert ids[r, : int(mask[r].sum())].tolist() == want, (r, text[:60])
        assert lengths[r] == len(want)
    enc = tok(["a great movie , really wonderful and moving from the first scene to the last", "bad"], padding="longest")
    assert enc["input_ids"][0][0] == 0 and enc["input_ids"][0][-1] == 2 and enc["input_ids"][1][-1] == 1
    assert enc["attention_mask"][1][-1] == 0
    short = tok(texts[:4], padding="max_length", max_length=16, return_tensors="np")
    assert short["input_ids"].shape == (4, 16) and (short["input_ids"][:, 0] == 0).all() and (short["input_ids"][:, -1] == 2).all()
    out = tmp_path / "saved"
    files = tok.save_pretrained(str(out))
    assert {os.path.basename(f) for f in files} == {"vocab.json", "merges.txt", "tokenizer_config.json", "special_tokens_map.json"}
    again = AutoTokenizer.from_pretrained(str(out))
    assert again(texts[:8], padding="longest") == tok(texts[:8], padding="longest")
    # byte-level BPE is lossless: decode(encode(x)) == x, including the odd ones
    for text in extra + texts[:5]:
        assert tok.decode(tok(text)["input_ids"]) == text


def test_text_data_source_is_reference_shaped():
    from huggingface_sagemaker_tensorflow_distributed_b200.data import build_datasets
    from huggingface_sagemaker_tensorflow_distributed_b200.models.configuration import resolve_config

    cfg = resolve_config("distilbert-base-uncased")
    tok = _tok(cfg.vocab_size)
    train, test = build_datasets(cfg, train_samples=40, test_samples=24, seq_len=512, seed=0, tokenizer=tok,
                                 data_source="text")
    for ds, n in ((train, 40), (test, 24)):
        assert ds.input_ids.shape == (n, 512) and ds.source == "synthetic-text"
        assert (ds.input_ids[:, 0] == 101).all() and ds.input_ids.max() < cfg.vocab_size
        assert ((ds.attention_mask == 0) <= (ds.input_ids == cfg.pad_token_id)).all()
        assert (ds.input_ids != tok.unk_token_id).all()  # the synthetic vocabulary covers its own text
        lens = ds.attention_mask.sum(1)
        assert lens.min() >= 8 and lens.max() <= 512 and len(set(lens.tolist())) > 5
    with pytest.raises(ValueError):
        build_datasets(cfg, train_samples=4, test_samples=4, seq_len=64, seed=0, tokenizer=None, data_source="text")


# ------------------------------------------------------------------------------------------ bucket manager
def _python_plan(offsets, ends, bucket_elems):
    spans, pb, cur_start, have = [], [], 0, False
    for o, e in zip(offsets, ends):
        if have and e - cur_start > bucket_elems:
            spans.append((cur_start, o))
            cur_start, have = o, False
        pb.append(len(spans))
        have = True
    spans.append((cur_start, ends[-1]))
    return spans, pb


def test_bucket_plan_matches_python_statement_and_model_layout():
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import host
    from huggingface_sagemaker_tensorflow_distributed_b200.parallel.dp_engine import plan_buckets

    rng = np.random.default_rng(0)
    for trial in range(20):
        n = int(rng.integers(1, 60))
        sizes = (rng.integers(1, 5000, size=n) + 7) // 8 * 8
        pads = rng.integers(0, 3, size=n) * 8
        offsets, ends, cur = [], [], 0
        for s, p in zip(sizes, pads):
            offsets.append(cur)
            ends.append(cur + int(s))
            cur += int(s + p)
        cap = int(rng.integers(64, 20000))
        assert host.bucket_plan(offsets, ends, cap) == _python_plan(offsets, ends, cap)
    # DistilBERT-sized arena through the DP engine's entry point: 32 MiB of bf16 per bucket
    slots, order, cur = {}, [], 0
    for i, numel in enumerate([768 * 2, 2, 768 * 768, 768] + [768 * 3072, 3072, 3072 * 768, 768, 2304 * 768, 2304] * 6 +
                              [30522 * 768]):
        name = f"p{i}"
        slots[name] = SimpleNamespace(offset=cur, end=cur + numel)
        order.append(name)
        cur += (numel + 63) // 64 * 64
    buckets = plan_buckets(slots, order, 32 * (1 << 20) // 2, world=8)
    assert [n for b in buckets for n in b.names] == order
    assert buckets[0].start == 0 and buckets[-1].end == slots[order[-1]].end
    assert all(a.end == b.start for a, b in zip(buckets, buckets[1:]))
    assert all(b.numel <= 32 * (1 << 20) // 2 or len(b.names) == 1 for b in buckets)
    with pytest.raises(Exception):
        host.bucket_plan([0, 4], [8, 12], 16)  # overlapping parameters


def test_embedding_block_is_one_row_sparse_bucket():
    """Real model layouts: encoder + head parameters are bucketed greedily, the embedding block (complete only
    when backward ends) is a single bucket that is a stack of hidden-size rows."""
    from huggingface_sagemaker_tensorflow_distributed_b200.models.configuration import resolve_config
    from huggingface_sagemaker_tensorflow_distributed_b200.models.engine import ALIGN, backward_ready_order, param_shapes
    from huggingface_sagemaker_tensorflow_distributed_b200.parallel.dp_engine import plan_buckets_with_embeddings
    import math

    for name in ("distilbert-base-uncased", "bert-base-uncased", "roberta-large"):
        cfg = resolve_config(name)
        shapes, slots, cur = param_shapes(cfg), {}, 0
        order = backward_ready_order(cfg)
        for n in order:
            numel = math.prod(shapes[n])
            slots[n] = SimpleNamespace(offset=cur, numel=numel, shape=shapes[n], end=cur + (numel + ALIGN - 1) // ALIGN * ALIGN)
            cur = slots[n].end
        for world in (2, 4, 8):
            bs = plan_buckets_with_embeddings(slots, order, 16 * (1 << 20) // 2, world, row_sparse=True)
            assert [n for b in bs for n in b.names] == order
            assert all(a.end == b.start for a, b in zip(bs, bs[1:])) and bs[-1].end == cur
            e = bs[-1]
            assert e.sparse and e.row_elems == cfg.hidden_size and e.vocab_rows == cfg.vocab_size
            assert e.names[-1] == "emb.word.weight" and all(n.startswith("emb.") for n in e.names)
            assert e.start + e.dense_rows * e.row_elems == slots["emb.word.weight"].offset
            assert e.numel % (8 * world) == 0 and not any(b.sparse for b in bs[:-1])
            assert len(bs) >= 6  # fine enough to overlap: 16 MiB buckets
        dense = plan_buckets_with_embeddings(slots, order, 16 * (1 << 20) // 2, 1, row_sparse=False)
        assert not dense[-1].sparse and dense[-1].names == bs[-1].names


def test_bucket_tracker_releases_in_order_and_honours_exemptions():
    from huggingface_sagemaker_tensorflow_distributed_b200.ops import host

    #            bucket:  0  0  0  1  1  2  3  3
    tr = host.BucketTracker([0, 0, 0, 1, 1, 2, 3, 3], [1, 0, 0, 0, 0, 0, 0, 0], 4)
    assert tr.mark([5]) == []            # bucket 2 complete, but 0 and 1 are not
    assert tr.mark([3, 4]) == []         # bucket 1 complete, still waiting for 0
    assert tr.mark([1]) == []
    assert tr.mark([2, 2]) == [0, 1, 2]  # param 0 is exempt; duplicates are harmless
    assert tr.done == 3
    assert tr.mark([7]) == [] and tr.mark([6]) == [3] and tr.done == 4
    tr.reset()
    assert tr.done == 0 and tr.mark([1, 2, 3, 4, 5, 6, 7]) == [0, 1, 2, 3]
    with pytest.raises(Exception):
        tr.mark([99])


# ------------------------------------------------------------------------------------------ loader
def test_native_loader_matches_python_batching():
    import torch

    from huggingface_sagemaker_tensorflow_distributed_b200.data import IGNORE_LABEL, BatchedDataset, synthetic_imdb
    from huggingface_sagemaker_tensorflow_distributed_b200.models.configuration import resolve_config

    cfg = resolve_config("tiny-bert")
    ds = synthetic_imdb(cfg, 53, 32)
    for rank, world in ((0, 1), (1, 2)):
        plain = BatchedDataset(ds, 8, rank, world)
        native = BatchedDataset(ds, 8, rank, world, prefetch=True, num_slots=3)
        assert len(plain) == len(native)
        for epoch in range(2):  # the ring is re-armed every epoch
            count = 0
            for (fa, la), (fb, lb) in zip(plain, native):
                assert torch.equal(fa["input_ids"], fb["input_ids"])
                assert torch.equal(fa["attention_mask"], fb["attention_mask"])
                assert torch.equal(la, lb)
                count += 1
            assert count == len(plain)
        assert (lb == IGNORE_LABEL).any()  # the short final batch was padded natively
    # early exit (steps_per_epoch) leaves the loader reusable
    capped = BatchedDataset(ds, 8, prefetch=True)
    it = iter(capped)
    first = next(it)[0]["input_ids"].clone()
    it.close()
    it = iter(capped)  # batches are views of the ring: valid while the iterator that produced them is alive
    assert torch.equal(next(it)[0]["input_ids"], first)
    it.close()
    assert len(list(BatchedDataset(ds, 8, prefetch=True, max_batches=2))) == 2
