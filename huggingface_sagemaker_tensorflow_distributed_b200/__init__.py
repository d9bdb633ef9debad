"""B200-native data-parallel fine-tuning of Hugging Face encoder classifiers — the
capabilities of philschmid/huggingface_sagemaker_tensorflow_distributed, rebuilt on
PyTorch + hand-written sm_100a CUDA kernels + NVLink-5 peer-to-peer collectives."""
__version__ = "0.1.0"

from .config import TrainConfig, parse_args  # noqa: F401
from .optim import Adam, AdamW, KerasAdamW, LRSchedule  # noqa: F401
from .trainer import (AutoModelForSequenceClassification, History, SequenceClassifier,  # noqa: F401
                      SparseCategoricalAccuracy, SparseCategoricalCrossentropy)
