"""Per-op device time of one engine training step (CUDA events, no profiler needed).
    python tools/op_profile.py [model] [batch] [seq]   -> gpurun_out/op_profile_<model>.txt"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from huggingface_sagemaker_tensorflow_distributed_b200 import trainer as T  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.data import BatchedDataset, synthetic_imdb  # noqa: E402
from huggingface_sagemaker_tensorflow_distributed_b200.optim import Adam  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "distilbert-base-uncased"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
S = int(sys.argv[3]) if len(sys.argv) > 3 else 512
model = T.AutoModelForSequenceClassification.from_pretrained(name, device=torch.device("cuda"), use_cuda_graph=False)
model.compile(optimizer=Adam(learning_rate=5e-5))
eng = model.backend
eng.setup_training(model.optimizer, 1000, B, S)
feats, labels = next(iter(BatchedDataset(synthetic_imdb(model.cfg, B, S, seed=1), B)))
eng.train_step(feats, labels)
rows = eng.profile(B, S)
tot = sum(r[2] for r in rows)
lines = [f"{name} B={B} S={S}: {tot:.0f} us/step over {sum(r[1] for r in rows)} ops (eager, event-timed; includes launch gaps)"]
for n, c, us in rows:
    lines.append(f"{us:9.1f} us {100 * us / tot:5.1f}%  x{c:<3d} {n}")
os.makedirs("gpurun_out", exist_ok=True)
open(f"gpurun_out/op_profile_{name}.txt", "w").write("\n".join(lines) + "\n")
print("\n".join(lines[:40]))
