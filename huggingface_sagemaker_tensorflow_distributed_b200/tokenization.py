"""Tokenizer front-end.

The reference does ``AutoTokenizer.from_pretrained(name)`` (scripts/train.py:69), calls the
tokenizer with ``truncation=True, padding=...`` inside ``dataset.map`` (scripts/train.py:75-77,
89-92), reads ``tokenizer.model_max_length`` (scripts/train.py:81) and finishes with
``tokenizer.save_pretrained(model_dir)`` (scripts/train.py:183).  ``WordPieceTokenizer`` gives
that surface on top of the native C++ tokenizer in libhstd_b200.so (csrc/tokenizer.cpp):

  * a directory with ``vocab.txt`` (any BERT / DistilBERT checkpoint) -> that vocabulary;
  * a hub name (no network on this image) -> a deterministic **synthetic** vocabulary of the
    model's size with the BERT special-token layout, so the text -> ids -> train -> save path is
    exercised end to end; ``source`` says which one was used.

``ByteLevelBPETokenizer`` is the same facade over the native byte-level BPE (csrc/bpe_tokenizer.cpp)
for a directory with ``vocab.json`` + ``merges.txt`` (RoBERTa checkpoints).  For a RoBERTa hub name
(no files, no network) ``AutoTokenizer`` serves the synthetic WordPiece vocabulary of the model's size.
"""
from __future__ import annotations

import json
import os
import tempfile
from typing import Dict, List, Optional, Sequence, Union

import numpy as np

SPECIAL_LAYOUT = {"[PAD]": 0, "[UNK]": 100, "[CLS]": 101, "[SEP]": 102, "[MASK]": 103}

_ONSETS = ["b", "br", "c", "ch", "cl", "d", "dr", "f", "fl", "g", "gr", "h", "j", "k", "l", "m", "n", "p", "pl", "pr",
           "qu", "r", "s", "sh", "sl", "st", "str", "t", "th", "tr", "v", "w", "wh", "y", "z", ""]
_NUCLEI = ["a", "e", "i", "o", "u", "ai", "ea", "ee", "ie", "oo", "ou", "oa"]
_CODAS = ["", "b", "ck", "d", "ft", "g", "l", "ll", "m", "n", "nd", "ng", "nt", "p", "r", "rd", "rk", "s", "sh", "ss",
          "st", "t", "th", "x"]
_SUFFIXES = ["s", "ed", "ing", "ly", "er", "est", "ness", "ment", "able", "ful", "less", "ish", "ive", "tion", "al", "y",
             "es", "en", "or", "ous"]
POSITIVE = ["great", "wonderful", "brilliant", "moving", "superb", "charming", "excellent", "delightful", "perfect",
            "masterpiece", "loved", "favorite", "beautiful", "funny", "touching", "fresh"]
NEGATIVE = ["awful", "boring", "terrible", "dull", "waste", "worst", "mess", "lazy", "bland", "painful", "hated",
            "forgettable", "ugly", "unfunny", "tedious", "stale"]
COMMON = ["the", "a", "and", "of", "to", "is", "it", "in", "this", "that", "was", "movie", "film", "i", "for", "with",
          "but", "as", "on", "not", "you", "are", "his", "her", "have", "be", "one", "all", "at", "by", "an", "they",
          "who", "so", "from", "like", "there", "or", "just", "about", "out", "if", "has", "what", "some", "good",
          "can", "more", "when", "very", "up", "no", "time", "even", "my", "would", "which", "story", "only", "really",
          "see", "their", "had", "were", "me", "well", "than", "we", "much", "been", "get", "will", "do", "also",
          "into", "people", "other", "first", "because", "how", "him", "most", "made", "its", "then", "way", "make",
          "them", "could", "too", "any", "after", "characters", "think", "watch", "two", "films", "character", "seen",
          "many", "being", "plot", "never", "acting", "little", "best", "where", "did", "ever", "know", "end", "scene"]


def _pseudo_words(n: int, rng: np.random.Generator) -> List[str]:
    seen, out = set(COMMON) | set(POSITIVE) | set(NEGATIVE), []
    while len(out) < n:
        m = 2 * (n - len(out)) + 64  # drawn in bulk: scalar Generator calls dominate otherwise
        syl = rng.integers(1, 4, size=m)
        o = rng.integers(len(_ONSETS), size=(m, 3))
        v = rng.integers(len(_NUCLEI), size=(m, 3))
        c = rng.integers(len(_CODAS), size=(m, 3))
        for i in range(m):
            w = "".join(_ONSETS[o[i, k]] + _NUCLEI[v[i, k]] + _CODAS[c[i, k]] for k in range(syl[i]))
            if 2 <= len(w) <= 14 and w not in seen:
                seen.add(w)
                out.append(w)
                if len(out) == n:
                    break
    return out


def build_synthetic_vocab(vocab_size: int, seed: int = 0) -> List[str]:
    """Deterministic WordPiece vocabulary with BERT's special-token ids ([PAD]=0, [UNK]=100,
    [CLS]=101, [SEP]=102, [MASK]=103, the rest of the first 999 slots ``[unusedN]``), every
    printable ASCII character as a word and as a ``##`` continuation (so nothing ASCII maps to
    [UNK]), the review lexicon, common suffix pieces and pseudo-words up to ``vocab_size``."""
    rng = np.random.default_rng(seed + 12345)
    toks: List[str] = []
    if vocab_size >= 2000:  # BERT layout
        reserved, by_id = 999, {v: k for k, v in SPECIAL_LAYOUT.items()}
    else:  # tiny test vocabularies: pack the specials at the front
        by_id = {i: k for i, k in enumerate(SPECIAL_LAYOUT)}
        reserved = len(by_id)
    unused = 0
    for i in range(reserved):
        if i in by_id:
            toks.append(by_id[i])
        else:
            toks.append(f"[unused{unused}]")
            unused += 1
    chars = [chr(c) for c in range(33, 127) if not ("A" <= chr(c) <= "Z")]
    body = chars + ["##" + c for c in chars if c.isalnum()]
    body += COMMON + POSITIVE + NEGATIVE + ["##" + s for s in _SUFFIXES]
    seen = set(toks)
    for t in body:
        if len(toks) >= vocab_size:
            break
        if t not in seen:
            seen.add(t)
            toks.append(t)
    remaining = vocab_size - len(toks)
    if remaining > 0:
        words = _pseudo_words(remaining, rng)
        # a quarter of the tail are continuation pieces, like a real WordPiece vocabulary
        for k, w in enumerate(words):
            cand = "##" + w if k % 4 == 3 else w
            if cand in seen:
                cand = w + str(k)
            seen.add(cand)
            toks.append(cand)
    return toks[:vocab_size]


def synthetic_reviews(num: int, seed: int = 0, split: str = "train", vocab: Optional[Sequence[str]] = None):
    """IMDB-shaped review *texts* with binary labels: log-normal word counts, common words plus
    vocabulary words (with random suffixes, capitalisation and punctuation so the pre-tokenizer
    has work to do) and a planted sentiment lexicon."""
    rng = np.random.default_rng(seed + (0 if split == "train" else 7919) + 31)
    words = [w for w in (vocab or []) if w.isalpha() and len(w) > 2][:5000] or COMMON
    labels = rng.integers(0, 2, size=num)
    n_words = np.clip(rng.lognormal(mean=5.2, sigma=0.6, size=num).astype(np.int64), 8, 900)
    texts = []
    for i in range(num):
        lex = POSITIVE if labels[i] else NEGATIVE
        n = int(n_words[i])
        kind = rng.random(n)
        pick_c = rng.integers(0, len(COMMON), size=n)
        pick_w = rng.integers(0, len(words), size=n)
        pick_l = rng.integers(0, len(lex), size=n)
        pick_s = rng.integers(0, len(_SUFFIXES), size=n)
        out, start = [], True
        for j in range(n):
            k = kind[j]
            if k < 0.55:
                w = COMMON[pick_c[j]]
            elif k < 0.63:
                w = lex[pick_l[j]]
            elif k < 0.90:
                w = words[pick_w[j]]
            else:
                w = words[pick_w[j]] + _SUFFIXES[pick_s[j]]
            if start:
                w = w.capitalize()
                start = False
            if k > 0.985:
                w += ","
            elif (j % 17) == 16 or j == n - 1:
                w += "!" if k > 0.5 and labels[i] else "."
                start = True
            out.append(w)
        texts.append(" ".join(out))
    return texts, labels.astype(np.int32)


class WordPieceTokenizer:
    """HF-tokenizer-shaped facade over the native WordPiece tokenizer.

    ``tok(texts, truncation=True, padding="longest" | "max_length" | False, max_length=None,
    return_tensors=None | "np" | "pt")`` returns ``{"input_ids", "attention_mask"}``."""

    vocab_files_names = {"vocab_file": "vocab.txt"}

    def __init__(self, vocab_file: str, do_lower_case: bool = True, strip_accents: Optional[bool] = None,
                 model_max_length: int = 512, source: str = "vocab.txt"):
        from .ops.host import NativeWordPiece

        self.vocab_file = str(vocab_file)
        self.do_lower_case = bool(do_lower_case)
        self.strip_accents = strip_accents
        self.model_max_length = int(model_max_length)
        self.source = source
        self._native = NativeWordPiece(self.vocab_file, self.do_lower_case, strip_accents)
        self.pad_token_id, self.unk_token_id = self._native.pad_id, self._native.unk_id
        self.cls_token_id, self.sep_token_id = self._native.cls_id, self._native.sep_id
        self.pad_token, self.unk_token, self.cls_token, self.sep_token, self.mask_token = (
            "[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]")

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_vocab(cls, tokens: Sequence[str], **kw) -> "WordPieceTokenizer":
        d = tempfile.mkdtemp(prefix="hstd_vocab_")
        path = os.path.join(d, "vocab.txt")
        with open(path, "w", encoding="utf-8") as fh:
            fh.write("\n".join(tokens) + "\n")
        return cls(path, **kw)

    @classmethod
    def synthetic(cls, vocab_size: int, model_max_length: int = 512, seed: int = 0) -> "WordPieceTokenizer":
        return cls.from_vocab(build_synthetic_vocab(vocab_size, seed), model_max_length=model_max_length,
                              source=f"synthetic(vocab_size={vocab_size}, seed={seed})")

    @classmethod
    def from_pretrained(cls, path: str, **kw) -> "WordPieceTokenizer":
        vocab = os.path.join(path, "vocab.txt")
        if not os.path.isfile(vocab):
            raise FileNotFoundError(f"{vocab} not found")
        cfg: Dict[str, object] = {}
        cfg_path = os.path.join(path, "tokenizer_config.json")
        if os.path.isfile(cfg_path):
            with open(cfg_path) as fh:
                cfg = json.load(fh)
        mml = cfg.get("model_max_length", 512)
        mml = 512 if not isinstance(mml, (int, float)) or mml > 1 << 20 else int(mml)
        return cls(vocab, do_lower_case=bool(kw.pop("do_lower_case", cfg.get("do_lower_case", True))),
                   strip_accents=kw.pop("strip_accents", cfg.get("strip_accents")),
                   model_max_length=int(kw.pop("model_max_length", mml)),
                   source=str(cfg.get("hstd_source", vocab)))

    def save_pretrained(self, save_dir: str) -> List[str]:
        """vocab.txt + tokenizer_config.json + special_tokens_map.json — the slow-tokenizer file set
        HF's BertTokenizer / AutoTokenizer reload (scripts/train.py:183)."""
        os.makedirs(save_dir, exist_ok=True)
        out = []
        vocab = os.path.join(save_dir, "vocab.txt")
        with open(vocab, "w", encoding="utf-8") as fh:
            for i in range(self.vocab_size):
                fh.write((self._native.id_to_token(i) or "") + "\n")
        out.append(vocab)
        cfg = {"tokenizer_class": "BertTokenizer", "do_lower_case": self.do_lower_case,
               "strip_accents": self.strip_accents, "model_max_length": self.model_max_length,
               "unk_token": self.unk_token, "sep_token": self.sep_token, "pad_token": self.pad_token,
               "cls_token": self.cls_token, "mask_token": self.mask_token, "tokenize_chinese_chars": True,
               "hstd_source": self.source}
        for name, payload in (("tokenizer_config.json", cfg),
                              ("special_tokens_map.json", {"unk_token": self.unk_token, "sep_token": self.sep_token,
                                                           "pad_token": self.pad_token, "cls_token": self.cls_token,
                                                           "mask_token": self.mask_token})):
            p = os.path.join(save_dir, name)
            with open(p, "w") as fh:
                json.dump(payload, fh, indent=2)
            out.append(p)
        return out

    # ------------------------------------------------------------------ queries
    @property
    def vocab_size(self) -> int:
        return self._native.vocab_size

    def __len__(self) -> int:
        return self.vocab_size

    def convert_tokens_to_ids(self, tokens: Union[str, Sequence[str]]):
        if isinstance(tokens, str):
            i = self._native.token_to_id(tokens)
            return self.unk_token_id if i < 0 else i
        return [self.convert_tokens_to_ids(t) for t in tokens]

    def convert_ids_to_tokens(self, ids: Union[int, Sequence[int]]):
        if isinstance(ids, (int, np.integer)):
            return self._native.id_to_token(int(ids))
        return [self._native.id_to_token(int(i)) for i in ids]

    def tokenize(self, text: str) -> List[str]:
        ids, mask, _ = self._native.encode_batch([text], max_len=max(8, 4 * len(text) + 2), add_special_tokens=False)
        return self.convert_ids_to_tokens(ids[0, : int(mask[0].sum())])

    def _special_ids(self) -> set:
        return {self.pad_token_id, self.cls_token_id, self.sep_token_id}

    def decode(self, ids: Sequence[int], skip_special_tokens: bool = True) -> str:
        """Inverse of the encoder up to the normalisation it applied (WordPiece: lower-casing, accent
        stripping, whitespace; continuation pieces are glued back, punctuation stays space-separated)."""
        toks = [self._native.id_to_token(int(i)) or self.unk_token for i in ids
                if not (skip_special_tokens and int(i) in self._special_ids())]
        out: List[str] = []
        for t in toks:
            if t.startswith("##") and out:
                out[-1] += t[2:]
            else:
                out.append(t)
        return " ".join(out)

    def batch_decode(self, batch: Sequence[Sequence[int]], skip_special_tokens: bool = True) -> List[str]:
        return [self.decode(r, skip_special_tokens) for r in batch]

    # ------------------------------------------------------------------ encoding
    def encode_dense(self, texts: Sequence[str], max_length: Optional[int] = None, padding: str = "max_length",
                     num_threads: int = 0):
        """int32 ``(ids, mask)`` arrays [n, L].  L = max_length for ``padding="max_length"``,
        the longest (truncated) row for ``"longest"``."""
        L = int(max_length or self.model_max_length)
        ids, mask, lengths = self._native.encode_batch(list(texts), L, True, num_threads)
        if padding == "longest" and len(texts):
            longest = int(min(L, max(2, int(lengths.max()))))
            ids, mask = np.ascontiguousarray(ids[:, :longest]), np.ascontiguousarray(mask[:, :longest])
        return ids, mask

    def __call__(self, text, truncation: bool = True, padding: Union[bool, str] = False,
                 max_length: Optional[int] = None, return_tensors: Optional[str] = None, **_unused):
        single = isinstance(text, str)
        texts = [text] if single else list(text)
        L = int(max_length or self.model_max_length)
        if not truncation:
            # the native encoder always bounds a row; "no truncation" = a bound no input reaches
            L = max(L, 2 + max((len(t) for t in texts), default=0))
        mode = "longest" if padding in (True, "longest") else "max_length" if padding == "max_length" else "none"
        ids, mask = self.encode_dense(texts, L, "longest" if mode != "max_length" else "max_length")
        if mode == "none":
            if return_tensors is not None and len(texts) > 1:
                raise ValueError("return_tensors needs padding when rows differ in length")
            rows = [ids[r, : int(mask[r].sum())].tolist() for r in range(len(texts))]
            masks = [[1] * len(r) for r in rows]
            if return_tensors is None:
                return {"input_ids": rows[0] if single else rows, "attention_mask": masks[0] if single else masks}
            ids, mask = np.asarray(rows, dtype=np.int32), np.asarray(masks, dtype=np.int32)
        if return_tensors == "np":
            return {"input_ids": ids, "attention_mask": mask}
        if return_tensors == "pt":
            import torch

            return {"input_ids": torch.from_numpy(ids.astype(np.int64)), "attention_mask": torch.from_numpy(mask.astype(np.int64))}
        if single:
            return {"input_ids": ids[0].tolist(), "attention_mask": mask[0].tolist()}
        return {"input_ids": ids.tolist(), "attention_mask": mask.tolist()}


_BPE_UNICODE_TO_BYTE: Optional[Dict[str, int]] = None


class ByteLevelBPETokenizer(WordPieceTokenizer):
    """RoBERTa / GPT-2 tokenizer files (``vocab.json`` + ``merges.txt``) served by the native byte-level
    BPE.  Inherits the calling conventions of ``WordPieceTokenizer``; <s> / </s> frame a sequence."""

    vocab_files_names = {"vocab_file": "vocab.json", "merges_file": "merges.txt"}

    def __init__(self, vocab_file: str, merges_file: str, model_max_length: int = 512, source: str = "vocab.json"):
        from .ops.host import NativeBPE

        self.vocab_file, self.merges_file = str(vocab_file), str(merges_file)
        self.do_lower_case, self.strip_accents = False, None
        self.model_max_length = int(model_max_length)
        self.source = source
        self._native = NativeBPE(self.vocab_file, self.merges_file)
        self.pad_token_id, self.unk_token_id = self._native.pad_id, self._native.unk_id
        self.cls_token_id, self.sep_token_id = self._native.cls_id, self._native.sep_id
        self.bos_token_id, self.eos_token_id = self.cls_token_id, self.sep_token_id
        self.pad_token, self.unk_token, self.cls_token, self.sep_token, self.mask_token = (
            "<pad>", "<unk>", "<s>", "</s>", "<mask>")

    def decode(self, ids: Sequence[int], skip_special_tokens: bool = True) -> str:
        """Exact inverse for byte-level BPE: tokens are concatenated and every code point is mapped
        back to its byte."""
        global _BPE_UNICODE_TO_BYTE
        if _BPE_UNICODE_TO_BYTE is None:
            keep = list(range(33, 127)) + list(range(161, 173)) + list(range(174, 256))
            table, extra = {}, 0
            for b in range(256):
                if b in keep:
                    table[chr(b)] = b
                else:
                    table[chr(256 + extra)] = b
                    extra += 1
            _BPE_UNICODE_TO_BYTE = table
        text = "".join(self._native.id_to_token(int(i)) or "" for i in ids
                       if not (skip_special_tokens and int(i) in self._special_ids()))
        return bytes(_BPE_UNICODE_TO_BYTE.get(ch, ord("?")) for ch in text).decode("utf-8", errors="replace")

    @classmethod
    def from_pretrained(cls, path: str, **kw) -> "ByteLevelBPETokenizer":
        vocab, merges = os.path.join(path, "vocab.json"), os.path.join(path, "merges.txt")
        if not (os.path.isfile(vocab) and os.path.isfile(merges)):
            raise FileNotFoundError(f"{vocab} / {merges} not found")
        cfg: Dict[str, object] = {}
        cfg_path = os.path.join(path, "tokenizer_config.json")
        if os.path.isfile(cfg_path):
            with open(cfg_path) as fh:
                cfg = json.load(fh)
        mml = cfg.get("model_max_length", 512)
        mml = 512 if not isinstance(mml, (int, float)) or mml > 1 << 20 else int(mml)
        return cls(vocab, merges, model_max_length=int(kw.pop("model_max_length", mml)), source=vocab)

    def save_pretrained(self, save_dir: str) -> List[str]:
        import shutil

        os.makedirs(save_dir, exist_ok=True)
        out = []
        for src, name in ((self.vocab_file, "vocab.json"), (self.merges_file, "merges.txt")):
            dst = os.path.join(save_dir, name)
            if os.path.abspath(src) != os.path.abspath(dst):
                shutil.copyfile(src, dst)
            out.append(dst)
        cfg = {"tokenizer_class": "RobertaTokenizer", "model_max_length": self.model_max_length, "add_prefix_space": False,
               "bos_token": "<s>", "eos_token": "</s>", "sep_token": "</s>", "cls_token": "<s>", "unk_token": "<unk>",
               "pad_token": "<pad>", "mask_token": "<mask>", "hstd_source": self.source}
        for name, payload in (("tokenizer_config.json", cfg),
                              ("special_tokens_map.json", {k: v for k, v in cfg.items() if k.endswith("_token")})):
            p = os.path.join(save_dir, name)
            with open(p, "w") as fh:
                json.dump(payload, fh, indent=2)
            out.append(p)
        return out


class AutoTokenizer:
    """``AutoTokenizer.from_pretrained`` of the reference (scripts/train.py:69)."""

    @staticmethod
    def from_pretrained(name_or_path: str, vocab_size: Optional[int] = None, model_max_length: int = 512,
                        seed: int = 0, **kw):
        if name_or_path and os.path.isdir(name_or_path):
            if os.path.isfile(os.path.join(name_or_path, "vocab.txt")):
                return WordPieceTokenizer.from_pretrained(name_or_path, **kw)
            if os.path.isfile(os.path.join(name_or_path, "vocab.json")) and os.path.isfile(os.path.join(name_or_path, "merges.txt")):
                return ByteLevelBPETokenizer.from_pretrained(name_or_path, **kw)
            try:  # byte-level BPE (RoBERTa) and anything else with local files: the HF implementation
                from transformers import AutoTokenizer as HFAuto

                return HFAuto.from_pretrained(name_or_path, local_files_only=True)
            except Exception:  # noqa: BLE001 - no tokenizer files in a weights-only directory
                pass
        if vocab_size is None:
            from .models.configuration import resolve_config

            mcfg = resolve_config(name_or_path)
            vocab_size, model_max_length = mcfg.vocab_size, mcfg.model_max_length
        return WordPieceTokenizer.synthetic(int(vocab_size), model_max_length=model_max_length, seed=seed)
