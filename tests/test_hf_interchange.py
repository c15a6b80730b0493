"""Hugging Face BERT checkpoints load into the sm_100a model and back (the reference's trials start from
``from_pretrained`` weights: examples/nlp/bert_glue_pytorch/model_def.py:60-68, bert_squad_pytorch/model_def.py)."""
import pytest
import torch

transformers = pytest.importorskip("transformers")

from determined_b200.models import bert as B  # noqa: E402


def _hf_config(num_labels=3):
    return transformers.BertConfig(vocab_size=500, hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=256,
                                   max_position_embeddings=64, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, num_labels=num_labels)


def _batch():
    torch.manual_seed(1)
    ids = torch.randint(0, 500, (3, 40))
    am = torch.ones(3, 40, dtype=torch.long)
    am[1, 30:] = 0
    tt = torch.randint(0, 2, (3, 40))
    return ids, am, tt


def test_sequence_classification_matches_transformers():
    torch.manual_seed(0)
    hf = transformers.BertForSequenceClassification(_hf_config()).eval()
    mine = B.BertForSequenceClassification(B.config_from_hf(hf.config)).eval()
    B.load_hf_state_dict(mine, hf.state_dict())
    ids, am, tt = _batch()
    y = torch.tensor([0, 2, 1])
    with torch.no_grad():
        ref = hf(input_ids=ids, attention_mask=am, token_type_ids=tt, labels=y)
        loss, logits = mine(ids, am, tt, y)
    assert torch.allclose(logits, ref.logits, atol=2e-4), (logits - ref.logits).abs().max()
    assert abs(loss.item() - ref.loss.item()) < 1e-4
    # and back: transformers loads what we export
    hf2 = transformers.BertForSequenceClassification(_hf_config()).eval()
    missing, unexpected = hf2.load_state_dict(B.to_hf_state_dict(mine), strict=False)
    assert not unexpected and all("position_ids" in k for k in missing)
    with torch.no_grad():
        assert torch.allclose(hf2(input_ids=ids, attention_mask=am, token_type_ids=tt).logits, ref.logits, atol=1e-5)


def test_question_answering_matches_transformers():
    torch.manual_seed(0)
    hf = transformers.BertForQuestionAnswering(_hf_config(2)).eval()
    mine = B.BertForQuestionAnswering(B.config_from_hf(hf.config)).eval()
    B.load_hf_state_dict(mine, hf.state_dict())
    ids, am, tt = _batch()
    st, en = torch.tensor([3, 10, 20]), torch.tensor([5, 12, 22])
    with torch.no_grad():
        ref = hf(input_ids=ids, attention_mask=am, token_type_ids=tt, start_positions=st, end_positions=en)
        loss, sl, el = mine(ids, am, tt, st, en)
    assert torch.allclose(sl, ref.start_logits, atol=2e-4) and torch.allclose(el, ref.end_logits, atol=2e-4)
    assert abs(loss.item() - ref.loss.item()) < 1e-4


def test_bare_encoder_checkpoint_and_mismatch():
    hf = transformers.BertModel(_hf_config())
    mine = B.BertForQuestionAnswering(B.config_from_hf(hf.config))
    B.load_hf_state_dict(mine, hf.state_dict())          # keys without the "bert." prefix; the span head keeps its init
    assert torch.equal(mine.layers[1].qkv_weight[128:256], hf.encoder.layer[1].attention.self.key.weight)
    small = B.BertForQuestionAnswering(B.BertConfig(vocab_size=500, hidden_size=128, num_hidden_layers=3, num_attention_heads=2,
                                                    intermediate_size=256, max_position_embeddings=64))
    with pytest.raises(KeyError):
        B.load_hf_state_dict(small, hf.state_dict())     # the checkpoint has no third layer
