"""SQuAD-shape trials on CPU: data layout, span metrics, and a tiny BERT / ALBERT span model learning through the
controller (reference counterparts: examples/nlp/bert_squad_pytorch, examples/nlp/albert_squad_pytorch)."""
import torch

from determined_b200.models import squad_trials as ST
from tests.harness_utils import TrainAndValidate, make_controller

TINY = {"global_batch_size": 4, "learning_rate": 2e-3, "num_hidden_layers": 1, "hidden_size": 64, "num_attention_heads": 1,
        "intermediate_size": 128, "vocab_size": 1200, "max_seq_length": 144, "dataset_size": 32, "validation_size": 8,
        "dropout": 0.0, "num_training_steps": 100, "max_query_length": 16, "max_answer_length": 6}


def test_synthetic_squad_layout():
    ds = ST.SyntheticSquadDataset(50, 160, 3000, 24, 10, seed=3)
    ids, am, tt, st, en = [torch.stack(t) for t in zip(*[ds[i] for i in range(50)])]
    assert ids.shape == (50, 160) and ((ids == 0) | (am == 1)).all()
    assert (st <= en).all() and (en - st < 10).all()
    assert am.gather(1, en[:, None]).all() and tt.gather(1, st[:, None]).all()      # the span lies inside the context segment


def test_span_metrics():
    sl = torch.full((3, 10), -5.0)
    el = torch.full((3, 10), -5.0)
    for i, (a, b) in enumerate([(2, 4), (1, 1), (6, 9)]):
        sl[i, a] = el[i, b] = 5.0
    m = ST.span_metrics(sl, el, torch.tensor([2, 3, 8]), torch.tensor([4, 5, 9]))
    assert abs(m["exact"].item() - 1 / 3) < 1e-6
    assert abs(m["f1"].item() - (1.0 + 0.0 + 2 * 0.5 * 1.0 / 1.5) / 3) < 1e-6


def _train(trial_cls, hp):
    tv = TrainAndValidate()
    make_controller(trial_cls, hp, tv.send(steps=2, validation_freq=2, scheduling_unit=8)).run()
    losses = [m["loss"] for m in tv.training_metrics]
    assert len(losses) == 16 and sum(losses[-4:]) < sum(losses[:4]), losses
    assert {"validation_loss", "exact", "f1"} <= set(tv.validation_metrics[-1])


def test_bert_squad_trial_learns():
    _train(ST.BertSquadTrial, TINY)


def test_albert_squad_trial_learns_with_shared_layer():
    hp = dict(TINY, num_hidden_layers=3, embedding_size=32)
    _train(ST.AlbertSquadTrial, hp)
    # the shared layer holds one parameter set regardless of depth
    from determined_b200.models.bert import AlbertConfig, AlbertForQuestionAnswering
    a = AlbertForQuestionAnswering(AlbertConfig(num_hidden_layers=2, hidden_size=64, num_attention_heads=1, intermediate_size=128, vocab_size=100))
    b = AlbertForQuestionAnswering(AlbertConfig(num_hidden_layers=9, hidden_size=64, num_attention_heads=1, intermediate_size=128, vocab_size=100))
    assert sum(p.numel() for p in a.parameters()) == sum(p.numel() for p in b.parameters())
