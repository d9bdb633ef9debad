"""Model math + checkpoint layout vs the transformers PyTorch classes (the oracle the
survey names, SURVEY.md §4 unit tier)."""
import json
import os

import pytest
import torch
import transformers

from huggingface_sagemaker_tensorflow_distributed_b200 import checkpoint
from huggingface_sagemaker_tensorflow_distributed_b200.models.configuration import KNOWN_MODELS, resolve_config
from huggingface_sagemaker_tensorflow_distributed_b200.models.hf_names import canonical_to_hf
from huggingface_sagemaker_tensorflow_distributed_b200.models.reference import EncoderForSequenceClassification

FAMILIES = [("tiny-distilbert", "DistilBertForSequenceClassification"), ("tiny-bert", "BertForSequenceClassification"),
            ("tiny-roberta", "RobertaForSequenceClassification")]


@pytest.mark.parametrize("name,hf_cls", FAMILIES)
def test_save_pretrained_reloads_into_transformers_with_same_logits(tmp_path, name, hf_cls):
    cfg = resolve_config(name)
    torch.manual_seed(0)
    m = EncoderForSequenceClassification(cfg).eval()
    checkpoint.save_pretrained(str(tmp_path), cfg, m.state_dict())
    assert sorted(os.listdir(tmp_path)) == ["config.json", "model.safetensors"]
    saved = json.load(open(tmp_path / "config.json"))
    assert saved["architectures"] == [hf_cls] and saved["model_type"] == cfg.model_type
    hm = getattr(transformers, hf_cls).from_pretrained(str(tmp_path)).eval()
    ids = torch.randint(3, cfg.vocab_size, (3, 24))
    mask = torch.ones_like(ids)
    mask[1, 15:] = 0
    ids[1, 15:] = cfg.pad_token_id
    with torch.no_grad():
        ours, theirs = m(ids, mask), hm(input_ids=ids, attention_mask=mask).logits
    assert torch.allclose(ours, theirs, atol=1e-5)
    # and back: --model_name_or_path <dir> is the reference's only form of resume
    cfg2 = resolve_config(str(tmp_path))
    assert cfg2.model_type == cfg.model_type and cfg2.hidden_size == cfg.hidden_size
    sd = checkpoint.load_pretrained_state(str(tmp_path), cfg2)
    for k, v in m.state_dict().items():
        assert torch.equal(sd[k], v), k


@pytest.mark.parametrize("name,params", [("distilbert-base-uncased", 66955010), ("bert-base-uncased", 109483778),
                                         ("roberta-base", 124647170), ("bert-large-uncased-whole-word-masking", 335143938),
                                         ("roberta-large", 355361794)])
def test_parameter_counts_match_survey_table(name, params):
    with torch.device("meta"):
        m = EncoderForSequenceClassification(resolve_config(name))
    assert m.num_parameters() == params


def test_hf_names_are_the_tf_reloadable_ones():
    cfg = resolve_config("tiny-distilbert")
    with torch.device("meta"):
        sd = EncoderForSequenceClassification(cfg).state_dict()
    names = set(canonical_to_hf(cfg, {k: torch.zeros(v.shape) for k, v in sd.items()}))
    assert "distilbert.transformer.layer.0.attention.q_lin.weight" in names
    assert "pre_classifier.weight" in names and "classifier.bias" in names


def test_unknown_model_is_a_clear_error():
    with pytest.raises(ValueError, match="no network"):
        resolve_config("some-org/some-model")
    assert "bert-large-uncased-whole-word-masking" in KNOWN_MODELS  # launch.py:17
