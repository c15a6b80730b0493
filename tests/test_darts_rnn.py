"""DARTS recurrent cell (reference: examples/hp_search_benchmarks/darts_penntreebank_pytorch): the regrouped schedule
against the step-by-step statement of the cell, the data pipeline, the averaged-SGD switch and the trial."""
import numpy as np
import torch

from determined_b200.models import darts_rnn as R
from determined_b200.models.darts_rnn_trial import BPTTBatchSampler, DARTSRNNTrial, PTBData, split_window, synthetic_corpus
from tests.harness_utils import TrainAndValidate, make_controller

GENO_HP = {"node1_edge": 0, "node2_edge": 1, "node3_edge": 1, "node4_edge": 1, "node5_edge": 2, "node6_edge": 5, "node7_edge": 3,
           "node8_edge": 5, "node1_op": "sigmoid", "node2_op": "relu", "node3_op": "relu", "node4_op": "identity",
           "node5_op": "tanh", "node6_op": "sigmoid", "node7_op": "tanh", "node8_op": "relu"}


def naive_cell(cell, inputs, hidden, x_mask=None, h_mask=None):
    """One concatenated GEMM for the initial state and one GEMM per node, per time step — the cell as published."""
    nh = cell.nhid
    h = hidden[0]
    outs = []
    for t in range(inputs.shape[0]):
        x = inputs[t] if x_mask is None else inputs[t] * x_mask
        hm = h if h_mask is None else h * h_mask
        c0, h0 = (torch.cat([x, hm], -1) @ cell._W0).split(nh, -1)
        states = [h + c0.sigmoid() * (R.ACTIVATIONS[cell.init_op](h0) - h)]
        for i, (name, pred) in enumerate(cell.genotype.recurrent):
            sp = states[pred]
            c, hh = ((sp if h_mask is None else sp * h_mask) @ cell._Ws[i]).split(nh, -1)
            states.append(sp + c.sigmoid() * (R.ACTIVATIONS[name](hh) - sp))
        h = torch.stack([states[i] for i in cell.genotype.concat], -1).mean(-1)
        outs.append(h)
    return torch.stack(outs)


def test_genotype_from_hparams_and_validation():
    g = R.genotype_from_hparams(GENO_HP)
    assert list(g.recurrent) == list(R.DARTS_V2.recurrent) and list(g.concat) == list(range(1, 9))
    bad = dict(GENO_HP, node2_edge=2)            # node 2 can read states 0..1 only
    try:
        R.genotype_from_hparams(bad)
        assert False
    except ValueError:
        pass


def test_regrouped_schedule_matches_the_published_cell():
    torch.manual_seed(0)
    cell = R.DARTSCell(24, 24, 0.25, 0.5, R.DARTS_V2).double()
    assert [p for p, _ in cell.groups] == [0, 1, 2, 3, 5] and dict(cell.groups)[1] == [1, 2, 3]     # 5 GEMMs instead of 8
    x, h0 = torch.randn(7, 3, 24, dtype=torch.double), torch.randn(1, 3, 24, dtype=torch.double) * 0.1
    cell.eval()
    out, last = cell(x, h0)
    ref = naive_cell(cell, x, h0)
    assert torch.allclose(out, ref, atol=1e-10) and torch.equal(last[0], out[-1])
    # training mode with given masks, and the gradients of every weight
    xm, hm = R.locked_mask(x, 3, 24, 0.5), R.locked_mask(x, 3, 24, 0.25)
    cell.train()
    out = cell(x, h0, masks=(xm, hm))[0]
    g1 = torch.autograd.grad(out.pow(2).sum(), list(cell.parameters()))
    ref = naive_cell(cell, x, h0, xm, hm)
    g2 = torch.autograd.grad(ref.pow(2).sum(), list(cell.parameters()))
    assert torch.allclose(out, ref, atol=1e-10)
    assert all(torch.allclose(a, b, atol=1e-8) for a, b in zip(g1, g2))


def test_model_ties_weights_and_dropouts_are_locked():
    torch.manual_seed(1)
    m = R.RNNModel(50, 16, 16, 16, dropout=0.5, dropouth=0.0, dropoutx=0.0, dropouti=0.0, dropoute=0.0)
    assert m.decoder.weight is m.encoder.weight
    tok = torch.randint(0, 50, (6, 4))
    m.train()
    logits, hidden, raw, dropped = m(tok, m.init_hidden(4))
    assert logits.shape == (6, 4, 50) and hidden[0].shape == (1, 4, 16)
    zero = dropped[0] == 0
    assert zero.any() and torch.equal(zero[0], zero[-1])           # the same units are dropped at every time step
    m.eval()
    a = m(tok, m.init_hidden(4))[0]
    b = m(tok, m.init_hidden(4))[0]
    assert torch.equal(a, b)


def test_bptt_sampler_windows_tile_the_stream():
    data = torch.arange(1000)
    ds = PTBData(data, 4)
    assert len(ds) == 250 and ds[1].tolist() == [1, 251, 501, 751]
    windows = list(BPTTBatchSampler(len(ds), 35, 20, seed=3))
    lens = [len(w) - 1 for w in windows]
    assert sum(lens) == 249 and max(lens) <= 55 and windows[0][0] == 0
    assert all(a[-1] == b[0] for a, b in zip(windows, windows[1:]))     # targets of one window end where the next begins
    assert len(set(lens)) > 2                                         # lengths vary
    x, y = split_window([ds[i] for i in windows[0]])
    assert torch.equal(x[1:], y[:-1]) and x.shape == (lens[0], 4)
    fixed = [len(w) - 1 for w in BPTTBatchSampler(len(ds), 35, 20, fixed=True)]
    assert set(fixed) == {35} and len(fixed) == 249 // 35
    valid = [len(w) - 1 for w in BPTTBatchSampler(len(ds), 35, 0, valid=True)]
    assert sum(valid) == 249 and set(valid[:-1]) == {35}


def test_synthetic_corpus_is_learnable_structure():
    c = synthetic_corpus(5000, 100, 1).numpy()
    follow = {}
    for a, b in zip(c[:-1], c[1:]):
        follow.setdefault(int(a), set()).add(int(b))
    assert np.mean([len(v) for v in follow.values()]) < 8          # few successors per word, not 100


def test_averaged_sgd_tracks_the_mean_of_the_iterates():
    p = torch.nn.Parameter(torch.zeros(3))
    av = R.AveragedSGD([p])
    av.after_step()
    assert av.ax == []
    av.set_optim("ASGD")
    seen = []
    for k in range(1, 5):
        p.data.fill_(float(k))
        av.after_step()
        seen.append(float(k))
        assert torch.allclose(av.ax[0], torch.full((3,), float(np.mean(seen))))
    av.swap()
    assert torch.allclose(p.data, torch.full((3,), 2.5)) and torch.allclose(av.ax[0], torch.full((3,), 4.0))
    av.swap()
    sd = av.state_dict()
    av2 = R.AveragedSGD([p])
    av2.load_state_dict(sd)
    assert av2.optim_name == "ASGD" and av2.steps == 4 and torch.equal(av2.ax[0], av.ax[0])


def test_darts_rnn_trial_learns_and_can_switch_to_averaging():
    hp = dict(GENO_HP, global_batch_size=8, learning_rate=20.0, emsize=32, nhid=32, nhidlast=32, bptt=12, max_seq_length_delta=6,
              dropout=0.1, dropouth=0.05, dropoutx=0.1, dropouti=0.05, dropoute=0.0, vocab_size=40, train_tokens=6000,
              valid_tokens=600, eval_batch_size=4, clip_gradients_l2_norm=0.25, beta=1e-3, alpha=0.0, weight_decay=8e-7,
              optimizer_switch_epoch=-1, nonmono=0)
    tv = TrainAndValidate()
    controller = make_controller(DARTSRNNTrial, hp, tv.send(steps=6, validation_freq=1, scheduling_unit=20))
    controller.run()
    ppl = [m["perplexity"] for m in tv.training_metrics]
    assert len(ppl) == 120 and np.mean(ppl[-10:]) < 0.5 * np.mean(ppl[:10]), (ppl[:3], ppl[-3:])
    assert {"loss", "perplexity"} <= set(tv.validation_metrics[-1])
    assert tv.validation_metrics[-1]["perplexity"] < 40 * 0.5
    trial = controller.trial
    assert len(trial._eval_history) == 6 and trial.averaging.optim_name == "SGD"      # validation kept improving: no switch
    assert trial._eval_history == sorted(trial._eval_history, reverse=True)           # the history holds the running best
    # a validation loss worse than the best seen `nonmono + 1` evaluations ago triggers the switch
    trial._last_loss = trial._eval_history[0] + 1.0
    trial._maybe_switch_optimizer()
    assert trial.averaging.optim_name == "ASGD" and len(trial.averaging.ax) == len(list(trial.model.parameters()))
    sd = controller.callbacks["averaging"].state_dict()
    assert set(sd) == {"averaging", "eval_history", "last_loss"}
