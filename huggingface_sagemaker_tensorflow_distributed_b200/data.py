"""Data pipeline.

The reference tokenises IMDB with the model's tokenizer, pads every example to
``tokenizer.model_max_length`` (512) and batches without shuffle / shard / repeat
(scripts/train.py:68-100).  There is no network, dataset cache or vocabulary on this image,
so the default source is a deterministic **synthetic IMDB-shaped** set: 25 000 train /
25 000 test examples, binary labels, review-like length distribution, padded to the maximum
sequence length, feeding only ``input_ids`` and ``attention_mask``.  A small planted signal
(label-dependent token frequencies) makes the loss curve meaningful.  When
``SM_CHANNEL_TRAIN`` / ``SM_CHANNEL_TEST`` point at pre-tokenised ``.npz``/``.pt`` files or
``.jsonl`` text plus a loadable tokenizer, those are used instead.  ``data_source="text"``
generates synthetic review *texts* and runs them through the native C++ WordPiece tokenizer
(csrc/tokenizer.cpp) — the reference's ``dataset.map(tokenize, batched=True)`` stage.
``data_source="hf:imdb"`` (or ``hf:<save_to_disk dir>``) is the reference's own ``load_dataset`` route, for
machines that have the dataset in their HF cache.

Batching: ``BatchedDataset`` assembles batches in Python; with ``prefetch=True`` a C++ worker
thread (csrc/batch_loader.cpp) stages them ahead of the training loop into a ring of pinned
slots, and a slot is recycled only after a CUDA event proves its H2D copies have run.

Sharding: by default rank r takes samples r, r+W, ... (DistributedSampler semantics) so the
job consumes each example once; ``no_shard=True`` restores the reference behaviour where
every rank iterates the full set.
"""
from __future__ import annotations

import glob
import json
import os
from dataclasses import dataclass
from typing import Dict, Iterator, Optional, Tuple

import numpy as np
import torch

from .models.configuration import EncoderConfig

IGNORE_LABEL = -1  # padding rows of a short final batch


@dataclass
class TokenizedDataset:
    input_ids: np.ndarray  # [N, S] int32
    attention_mask: np.ndarray  # [N, S] int32 (1 = token, 0 = pad)
    labels: np.ndarray  # [N] int32
    source: str = "synthetic"

    def __len__(self) -> int:
        return int(self.labels.shape[0])

    @property
    def seq_len(self) -> int:
        return int(self.input_ids.shape[1])


def synthetic_imdb(cfg: EncoderConfig, num_samples: int, seq_len: int, seed: int = 0,
                   split: str = "train") -> TokenizedDataset:
    """IMDB-shaped synthetic token data.  Lengths follow a log-normal fitted to IMDB
    word-piece counts (median ≈ 230, long tail truncated at ``seq_len``)."""
    rng = np.random.default_rng(seed + (0 if split == "train" else 7919))
    lengths = np.clip(rng.lognormal(mean=5.45, sigma=0.62, size=num_samples).astype(np.int64), 16, seq_len)
    labels = rng.integers(0, 2, size=num_samples, dtype=np.int64)
    special = 4 if cfg.model_type == "roberta" else 1000  # keep clear of [PAD]/[CLS]/[SEP]/unused ids
    lo = min(special, cfg.vocab_size // 2)
    ids = rng.integers(lo, cfg.vocab_size, size=(num_samples, seq_len), dtype=np.int64)
    # planted sentiment: ~6% of tokens come from a label-specific 64-token lexicon
    lex = lo + np.arange(128) % max(1, cfg.vocab_size - lo)
    plant = rng.random((num_samples, seq_len)) < 0.06
    which = rng.integers(0, 64, size=(num_samples, seq_len))
    planted = lex[(which + 64 * labels[:, None]) % 128]
    ids = np.where(plant, planted, ids)
    pos = np.arange(seq_len)[None, :]
    mask = (pos < lengths[:, None]).astype(np.int32)
    cls_id, sep_id = (0, 2) if cfg.model_type == "roberta" else (min(101, cfg.vocab_size - 2), min(102, cfg.vocab_size - 1))
    ids[:, 0] = cls_id
    ids[np.arange(num_samples), lengths - 1] = sep_id
    ids = np.where(mask == 1, ids, cfg.pad_token_id).astype(np.int32)
    return TokenizedDataset(ids, mask, labels.astype(np.int32), source="synthetic")


def synthetic_imdb_text(cfg: EncoderConfig, tokenizer, num_samples: int, seq_len: int, seed: int = 0,
                        split: str = "train", padding: str = "max_length") -> TokenizedDataset:
    """Synthetic review texts -> native tokenizer -> dense ``[N, seq_len]`` tables: steps (1)-(5) of
    the reference pipeline (scripts/train.py:69-83).  ``padding="longest"`` tokenises like the
    reference's train split and is then densified to ``seq_len`` with the pad id, as
    ``to_tensor(shape=[None, model_max_length])`` does (scripts/train.py:80-83)."""
    from .tokenization import synthetic_reviews

    vocab = [tokenizer.convert_ids_to_tokens(i) for i in range(min(tokenizer.vocab_size, 12000))]
    texts, labels = synthetic_reviews(num_samples, seed, split, vocab)
    if hasattr(tokenizer, "encode_dense"):
        ids, mask = tokenizer.encode_dense(texts, max_length=seq_len, padding=padding)
    else:  # an HF tokenizer (e.g. a local RoBERTa BPE)
        enc = tokenizer(texts, truncation=True, padding=padding, max_length=seq_len)
        ids, mask = np.asarray(enc["input_ids"], dtype=np.int32), np.asarray(enc["attention_mask"], dtype=np.int32)
    ids, mask = _pad_to(ids, seq_len, 0), _pad_to(mask, seq_len, 0)
    ids = np.where(mask == 1, ids, cfg.pad_token_id).astype(np.int32)
    if int(ids.max()) >= cfg.vocab_size:
        raise ValueError(f"tokenizer produced id {int(ids.max())} >= model vocab_size {cfg.vocab_size}")
    return TokenizedDataset(np.ascontiguousarray(ids), np.ascontiguousarray(mask.astype(np.int32)), labels,
                            source="synthetic-text")


def _pad_to(a: np.ndarray, seq_len: int, value: int) -> np.ndarray:
    if a.shape[1] == seq_len:
        return a
    if a.shape[1] > seq_len:
        return a[:, :seq_len]
    out = np.full((a.shape[0], seq_len), value, dtype=a.dtype)
    out[:, :a.shape[1]] = a
    return out


def load_channel(path: str, cfg: EncoderConfig, seq_len: int, tokenizer=None) -> Optional[TokenizedDataset]:
    """Reads a SageMaker input channel directory: ``*.npz`` / ``*.pt`` with input_ids,
    attention_mask, label(s); or ``*.jsonl`` with text + label (needs a tokenizer)."""
    if not path or not os.path.isdir(path):
        return None
    for f in sorted(glob.glob(os.path.join(path, "*.npz"))):
        z = np.load(f)
        lab = z["labels"] if "labels" in z else z["label"]
        return TokenizedDataset(_pad_to(z["input_ids"].astype(np.int32), seq_len, cfg.pad_token_id),
                                _pad_to(z["attention_mask"].astype(np.int32), seq_len, 0), lab.astype(np.int32), f)
    for f in sorted(glob.glob(os.path.join(path, "*.pt"))):
        z = torch.load(f, map_location="cpu", weights_only=True)
        lab = z["labels"] if "labels" in z else z["label"]
        return TokenizedDataset(_pad_to(z["input_ids"].numpy().astype(np.int32), seq_len, cfg.pad_token_id),
                                _pad_to(z["attention_mask"].numpy().astype(np.int32), seq_len, 0),
                                lab.numpy().astype(np.int32), f)
    files = sorted(glob.glob(os.path.join(path, "*.jsonl")))
    if files and tokenizer is not None:
        texts, labels = [], []
        for f in files:
            with open(f) as fh:
                for line in fh:
                    if line.strip():
                        r = json.loads(line)
                        texts.append(r["text"])
                        labels.append(int(r["label"]))
        enc = tokenizer(texts, truncation=True, padding="max_length", max_length=seq_len)
        return TokenizedDataset(np.asarray(enc["input_ids"], dtype=np.int32),
                                np.asarray(enc["attention_mask"], dtype=np.int32),
                                np.asarray(labels, dtype=np.int32), files[0])
    return None


def load_hf_text_dataset(cfg: EncoderConfig, tokenizer, seq_len: int, name_or_dir: str = "imdb",
                         max_train: int = -1, max_test: int = -1) -> Optional[Tuple[TokenizedDataset, TokenizedDataset]]:
    """The reference's real-data path (scripts/train.py:69-100): ``load_dataset("imdb", split=["train","test"])``
    -> ``tokenizer(text, truncation=True, padding=...)`` -> dense ``[N, model_max_length]`` ids / mask + label.
    ``name_or_dir`` is a hub dataset name served from the local HF cache (there is no network here:
    ``HF_DATASETS_OFFLINE``) or a directory written by ``DatasetDict.save_to_disk``.  Returns None when the
    ``datasets`` library, the data or a tokenizer is not available, so callers fall back to synthetic data."""
    if tokenizer is None:
        return None
    try:
        import datasets as hfds

        if os.path.isdir(name_or_dir):
            dd = hfds.load_from_disk(name_or_dir)
            train, test = dd["train"], dd["test"]
        else:
            train, test = hfds.load_dataset(name_or_dir, split=["train", "test"])
    except Exception:  # noqa: BLE001 - not cached / offline / library missing
        return None
    out = []
    for split, padding, cap in ((train, "longest", max_train), (test, "max_length", max_test)):
        if cap is not None and cap >= 0:
            split = split.select(range(min(cap, len(split))))
        texts, labels = list(split["text"]), np.asarray(split["label"], dtype=np.int32)
        if hasattr(tokenizer, "encode_dense"):  # the native tokenizers
            ids, mask = tokenizer.encode_dense(texts, max_length=seq_len, padding=padding)
        else:
            enc = tokenizer(texts, truncation=True, padding=padding, max_length=seq_len)
            ids, mask = np.asarray(enc["input_ids"], dtype=np.int32), np.asarray(enc["attention_mask"], dtype=np.int32)
        ids, mask = _pad_to(np.asarray(ids, dtype=np.int32), seq_len, 0), _pad_to(np.asarray(mask, dtype=np.int32), seq_len, 0)
        ids = np.where(mask == 1, ids, cfg.pad_token_id).astype(np.int32)
        if int(ids.max()) >= cfg.vocab_size:
            raise ValueError(f"tokenizer produced id {int(ids.max())} >= model vocab_size {cfg.vocab_size}")
        out.append(TokenizedDataset(np.ascontiguousarray(ids), np.ascontiguousarray(mask), labels, source=f"hf:{name_or_dir}"))
    return out[0], out[1]


def try_load_tokenizer(name_or_path: Optional[str]):
    """``AutoTokenizer.from_pretrained`` (scripts/train.py:69): the tokenizer files of a local
    directory when present, otherwise a synthetic WordPiece vocabulary of the model's size served
    by the native tokenizer.  None only if the native library has not been built."""
    if not name_or_path:
        return None
    try:
        from .tokenization import AutoTokenizer

        return AutoTokenizer.from_pretrained(name_or_path)
    except Exception:  # noqa: BLE001 - e.g. libhstd_b200.so absent on a source-only checkout
        return None


class BatchedDataset:
    """``tf.data.Dataset.from_tensor_slices(...).batch(bs)`` stand-in: a re-iterable of
    ``(features, labels)`` batches with static shapes.  A short final batch is padded with
    rows whose label is ``IGNORE_LABEL`` (they contribute to neither loss nor accuracy), so
    kernels and CUDA graphs see one shape."""

    def __init__(self, ds: TokenizedDataset, batch_size: int, rank: int = 0, world_size: int = 1,
                 shard: bool = True, max_batches: int = -1, pin_memory: bool = False, prefetch: bool = False,
                 num_slots: int = 4):
        self.ds, self.batch_size = ds, int(batch_size)
        self.prefetch, self.num_slots = bool(prefetch), max(3, int(num_slots))
        self._loader = None
        idx = np.arange(len(ds))
        # Every rank must take the same number of steps: the fused all-reduce is a collective, and a rank
        # that runs one batch more than its peers would wait for them in the flag barrier.  With
        # len(ds) % world != 0 the strided shards differ by one sample, so the batch count is derived
        # from the LONGEST shard (DistributedSampler semantics); a rank that runs out of samples pads
        # with rows labelled IGNORE_LABEL, which count towards neither loss nor accuracy.
        self._common_len = len(idx)
        if shard and world_size > 1:
            self._common_len = (len(idx) + world_size - 1) // world_size
            idx = idx[rank::world_size]
        self.indices = idx
        self.max_batches = max_batches
        self.pin_memory = pin_memory and torch.cuda.is_available()

    def __len__(self) -> int:
        n = (self._common_len + self.batch_size - 1) // self.batch_size
        return n if self.max_batches < 0 else min(n, self.max_batches)

    @property
    def num_samples(self) -> int:
        return min(len(self.indices), len(self) * self.batch_size)

    # ------------------------------------------------------------------ native prefetching path
    def _ensure_loader(self):
        if self._loader is not None:
            return self._loader
        from .ops.host import NativeLoader

        bs, S = self.batch_size, self.ds.seq_len
        mk = (lambda *shape: torch.empty(*shape, dtype=torch.int32).pin_memory()) if self.pin_memory else \
             (lambda *shape: torch.empty(*shape, dtype=torch.int32))
        self._slots = [(mk(bs, S), mk(bs, S), mk(bs)) for _ in range(self.num_slots)]
        tables = tuple(np.ascontiguousarray(a, dtype=np.int32) for a in
                       (self.ds.input_ids, self.ds.attention_mask, self.ds.labels))
        self._loader = NativeLoader(*tables, self.indices, bs, len(self),
                                    [tuple(t.data_ptr() for t in slot) for slot in self._slots],
                                    pad_id=0, ignore_label=IGNORE_LABEL)
        return self._loader

    def _iter_prefetch(self) -> Iterator[Tuple[Dict[str, torch.Tensor], torch.Tensor]]:
        """Yields views of the pinned ring.  A batch stays valid until two more batches have been
        requested; its slot goes back to the worker only after an event recorded behind the
        consumer's work on that batch has completed, so the host also never runs more than two
        steps ahead of the device."""
        ld = self._ensure_loader()
        ld.start_epoch()
        cuda = self.pin_memory
        held: list = []  # [(slot, event-or-None)] oldest first
        try:
            while True:
                if held and cuda and held[-1][1] is None:
                    ev = torch.cuda.Event()
                    ev.record()
                    held[-1] = (held[-1][0], ev)
                while len(held) >= 2:
                    slot, ev = held.pop(0)
                    if ev is not None:
                        ev.synchronize()
                    ld.release(slot)
                slot = ld.next()
                if slot < 0:
                    break
                held.append((slot, None))
                ids, mask, lab = self._slots[slot]
                yield {"input_ids": ids, "attention_mask": mask}, lab
        finally:
            if cuda:
                torch.cuda.synchronize()
            for slot, _ in held:
                ld.release(slot)

    def __iter__(self) -> Iterator[Tuple[Dict[str, torch.Tensor], torch.Tensor]]:
        if self.prefetch:
            yield from self._iter_prefetch()
            return
        bs, S = self.batch_size, self.ds.seq_len
        for b in range(len(self)):
            sel = self.indices[b * bs:(b + 1) * bs]
            ids = np.full((bs, S), 0, dtype=np.int32)
            mask = np.zeros((bs, S), dtype=np.int32)
            lab = np.full((bs,), IGNORE_LABEL, dtype=np.int32)
            ids[:len(sel)] = self.ds.input_ids[sel]
            mask[:len(sel)] = self.ds.attention_mask[sel]
            lab[:len(sel)] = self.ds.labels[sel]
            if len(sel) < bs:  # keep padded rows well-formed: one real token
                mask[len(sel):, 0] = 1
            feats = {"input_ids": torch.from_numpy(ids), "attention_mask": torch.from_numpy(mask)}
            labels = torch.from_numpy(lab)
            if self.pin_memory:
                feats = {k: v.pin_memory() for k, v in feats.items()}
                labels = labels.pin_memory()
            yield feats, labels


def build_datasets(cfg: EncoderConfig, *, train_samples: int, test_samples: int, seq_len: int, seed: int,
                   train_channel: Optional[str] = None, test_channel: Optional[str] = None,
                   tokenizer=None, data_source: str = "ids") -> Tuple[TokenizedDataset, TokenizedDataset]:
    """``data_source``: "ids" draws synthetic token ids directly; "text" generates review texts and
    tokenises them (needs ``tokenizer``).  Channel directories, when given and readable, win."""
    train = load_channel(train_channel, cfg, seq_len, tokenizer) if train_channel else None
    test = load_channel(test_channel, cfg, seq_len, tokenizer) if test_channel else None
    if data_source == "text" and tokenizer is None:
        raise ValueError("data_source='text' needs a tokenizer")
    if data_source.startswith("hf:") and (train is None or test is None):
        # hf:imdb (local HF cache) or hf:/path/to/save_to_disk dir — the reference's load_dataset route
        got = load_hf_text_dataset(cfg, tokenizer, seq_len, data_source[3:], train_samples, test_samples)
        if got is None:
            raise ValueError(f"data_source={data_source!r}: dataset, `datasets` library or tokenizer unavailable offline")
        train, test = (train or got[0]), (test or got[1])
        data_source = "ids"
    if train is None:
        # the reference pads the train split to the longest example and the test split to max_length
        # before densifying both to model_max_length (scripts/train.py:76,90)
        train = (synthetic_imdb_text(cfg, tokenizer, train_samples, seq_len, seed, "train", padding="longest")
                 if data_source == "text" else synthetic_imdb(cfg, train_samples, seq_len, seed, "train"))
    if test is None:
        test = (synthetic_imdb_text(cfg, tokenizer, test_samples, seq_len, seed, "test", padding="max_length")
                if data_source == "text" else synthetic_imdb(cfg, test_samples, seq_len, seed, "test"))
    return train, test
