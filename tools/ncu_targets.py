"""One eager training step of DistilBERT (B=8,S=512) for ncu kernel captures."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from huggingface_sagemaker_tensorflow_distributed_b200 import trainer as T
from huggingface_sagemaker_tensorflow_distributed_b200.data import BatchedDataset, synthetic_imdb
from huggingface_sagemaker_tensorflow_distributed_b200.optim import Adam
model = T.AutoModelForSequenceClassification.from_pretrained("distilbert-base-uncased", device=torch.device("cuda"), use_cuda_graph=False)
model.compile(optimizer=Adam(learning_rate=5e-5))
eng = model.backend
eng.setup_training(model.optimizer, 1000, 8, 512)
feats, labels = next(iter(BatchedDataset(synthetic_imdb(model.cfg, 8, 512, seed=1), 8)))
for _ in range(int(os.environ.get("STEPS", "2"))):
    eng.train_step(feats, labels)
torch.cuda.synchronize()
print("ok")
