"""Host-side throughput of the native tokenizers vs the HF `tokenizers` (Rust) library on the same
vocabulary and the same synthetic IMDB-shaped reviews (CPU only)."""
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from huggingface_sagemaker_tensorflow_distributed_b200.tokenization import (ByteLevelBPETokenizer, WordPieceTokenizer,  # noqa: E402
                                                                             synthetic_reviews)

N = int(os.environ.get("N", "5000"))
tok = WordPieceTokenizer.synthetic(30522)
vocab = [tok.convert_ids_to_tokens(i) for i in range(1000, 9000)]
texts, _ = synthetic_reviews(N, seed=0, vocab=vocab)
words = sum(len(t.split()) for t in texts)
print(f"{N} reviews, {words} words, {os.cpu_count()} CPUs")


def best(fn, reps=3):
    t = 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        t = min(t, time.perf_counter() - t0)
    return t


t_native = best(lambda: tok.encode_dense(texts, 512))
print(f"WordPiece  native C++ : {N / t_native:9.0f} reviews/s")
try:
    import tokenizers

    d = tempfile.mkdtemp()
    tok.save_pretrained(d)
    hf = tokenizers.BertWordPieceTokenizer(os.path.join(d, "vocab.txt"), lowercase=True)
    hf.enable_truncation(512)
    hf.enable_padding(length=512)
    t_hf = best(lambda: hf.encode_batch(texts))
    print(f"WordPiece  HF tokenizers: {N / t_hf:9.0f} reviews/s   (native / HF = {t_hf / t_native:.2f}x)")
    o = tokenizers.ByteLevelBPETokenizer()
    o.train_from_iterator(texts[:2000], vocab_size=8000, special_tokens=["<s>", "<pad>", "</s>", "<unk>", "<mask>"], show_progress=False)
    o.save_model(d)
    bpe = ByteLevelBPETokenizer.from_pretrained(d)
    t_nb = best(lambda: bpe.encode_dense(texts, 512))
    o2 = tokenizers.ByteLevelBPETokenizer(os.path.join(d, "vocab.json"), os.path.join(d, "merges.txt"))
    o2.enable_truncation(512)
    o2.enable_padding(length=512)
    t_hb = best(lambda: o2.encode_batch(texts))
    print(f"BPE        native C++ : {N / t_nb:9.0f} reviews/s")
    print(f"BPE        HF tokenizers: {N / t_hb:9.0f} reviews/s   (native / HF = {t_hb / t_nb:.2f}x)")
except ImportError:
    print("tokenizers library not installed: no comparison")
