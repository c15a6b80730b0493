"""``DARTSRNNTrial`` — a DARTS recurrent-cell language model inside ``PyTorchTrial``; the architecture is given by
``node{i}_edge`` / ``node{i}_op`` hyperparameters so that a searcher (``adaptive_asha`` in the reference's
``adaptive.yaml``) explores the space.

Counterpart of the reference's ``examples/hp_search_benchmarks/darts_penntreebank_pytorch/model_def.py:56-303`` and
``data.py:11-88``: variable-length BPTT windows (``N(bptt, 5)``, occasionally ``bptt / 2``, capped at ``bptt +
max_seq_length_delta``) with the learning rate scaled by ``seq_len / bptt``, the hidden state carried from batch to batch
(detached), activation regularisers ``alpha`` / ``beta``, gradient clipping at ``clip_gradients_l2_norm``, a switch from
SGD to averaged SGD when the validation loss has not improved for ``nonmono`` evaluations after ``optimizer_switch_epoch``,
``evaluate_full_dataset`` with the averaged weights, and ``loss / raw_loss / perplexity`` metrics.

Data: ``data.data_download_dir`` is read when it holds ``train.txt`` / ``valid.txt`` (the Penn Treebank files of the
reference); there is no network here, so otherwise a synthetic corpus of the same shape is generated — a 10,000-word
vocabulary driven by a sparse random bigram table, which a model can learn (perplexity falls well below the uniform
10,000).
"""
from __future__ import annotations

import logging
import math
import os
from typing import Any, Dict, Iterator, List, Optional

import numpy as np
import torch
from torch import nn

from .. import pytorch as dp
from .darts_rnn import AveragedSGD, RNNModel, genotype_from_hparams


# ---- corpus -------------------------------------------------------------------------------------------------------------------
def synthetic_corpus(n_tokens: int, vocab: int, seed: int, branching: int = 4) -> torch.Tensor:
    """Tokens from a first-order chain in which every word has ``branching`` likely successors."""
    rng = np.random.RandomState(seed)
    table = np.random.RandomState(12345).randint(0, vocab, size=(vocab, branching))   # shared by train and validation
    out = np.empty(n_tokens, dtype=np.int64)
    cur = int(rng.randint(vocab))
    choice = rng.randint(0, branching, size=n_tokens)
    jump = rng.rand(n_tokens) < 0.05
    jump_to = rng.randint(0, vocab, size=n_tokens)
    for i in range(n_tokens):
        cur = int(jump_to[i]) if jump[i] else int(table[cur, choice[i]])
        out[i] = cur
    return torch.from_numpy(out)


class Corpus:
    """``train`` / ``valid`` token tensors and the vocabulary size, from text files when present."""

    def __init__(self, path: Optional[str], vocab: int, train_tokens: int, valid_tokens: int) -> None:
        if path and os.path.exists(os.path.join(path, "train.txt")) and os.path.exists(os.path.join(path, "valid.txt")):
            self.word2idx: Dict[str, int] = {}
            self.train = self._tokenize(os.path.join(path, "train.txt"))
            self.valid = self._tokenize(os.path.join(path, "valid.txt"))
            self.ntokens = len(self.word2idx)
        else:
            self.train = synthetic_corpus(train_tokens, vocab, 1)
            self.valid = synthetic_corpus(valid_tokens, vocab, 2)
            self.ntokens = vocab

    def _tokenize(self, path: str) -> torch.Tensor:
        ids: List[int] = []
        with open(path, "r", encoding="utf8") as f:
            for line in f:
                for word in line.split() + ["<eos>"]:
                    ids.append(self.word2idx.setdefault(word, len(self.word2idx)))
        return torch.tensor(ids, dtype=torch.int64)


class PTBData(torch.utils.data.Dataset):
    """The token stream folded into ``batch_size`` parallel columns; item ``i`` is row ``i`` ([batch_size] tokens)."""

    def __init__(self, data: torch.Tensor, batch_size: int) -> None:
        n = data.size(0) // batch_size
        self.data = data[:n * batch_size].view(batch_size, n).t().contiguous()

    def __len__(self) -> int:
        return len(self.data)

    def __getitem__(self, i: int) -> torch.Tensor:
        return self.data[i]


class BPTTBatchSampler(torch.utils.data.Sampler):
    """Consecutive windows of rows: ``seq_len + 1`` indices each (inputs are the first ``seq_len`` rows, targets the rows
    shifted by one); the window length is random in training and ``bptt`` in validation or when ``fixed`` is set."""

    def __init__(self, n_rows: int, bptt: int, max_delta: int, valid: bool = False, fixed: bool = False, seed: int = 0) -> None:
        self.n_rows, self.bptt, self.max_len, self.static = n_rows - 1, bptt, bptt + max_delta, valid or fixed
        self.drop_tail = fixed
        self.rng = np.random.RandomState(seed)

    def _seq_len(self, i: int) -> int:
        if self.static:
            n = self.bptt
        else:
            base = self.bptt if self.rng.random_sample() < 0.95 else self.bptt / 2.0
            n = min(max(5, int(self.rng.normal(base, 5))), self.max_len)
        return min(n, self.n_rows - i)

    def __len__(self) -> int:
        return max(self.n_rows // self.bptt, 1)

    def __iter__(self) -> Iterator[List[int]]:
        i = 0
        while i < self.n_rows:
            n = self._seq_len(i)
            if n <= 0 or (self.drop_tail and n < self.bptt):
                break
            yield list(range(i, i + n + 1))
            i += n


def split_window(rows: List[torch.Tensor]):
    w = torch.stack(rows)
    return w[:-1], w[1:]


# ---- trial --------------------------------------------------------------------------------------------------------------------
class _AveragingState(dp.PyTorchCallback):
    """Checkpoints the averaged-SGD state and the evaluation history with the trial."""

    def __init__(self, trial: "DARTSRNNTrial") -> None:
        self.trial = trial

    def state_dict(self) -> Dict[str, Any]:
        return {"averaging": self.trial.averaging.state_dict(), "eval_history": list(self.trial._eval_history),
                "last_loss": self.trial._last_loss}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        self.trial.averaging.load_state_dict(sd["averaging"])
        self.trial._eval_history, self.trial._last_loss = list(sd["eval_history"]), sd["last_loss"]


class DARTSRNNTrial(dp.PyTorchTrial):
    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = dict(context.get_hparams())
        self.hp = hp
        data_cfg = context.get_data_config() or {}
        self.corpus = Corpus(data_cfg.get("data_download_dir"), int(hp.get("vocab_size", 10000)),
                             int(hp.get("train_tokens", 200_000)), int(hp.get("valid_tokens", 20_000)))
        size = int(hp.get("emsize", 850))
        model = RNNModel(self.corpus.ntokens, size, int(hp.get("nhid", size)), int(hp.get("nhidlast", size)),
                         float(hp.get("dropout", 0.75)), float(hp.get("dropouth", 0.25)), float(hp.get("dropoutx", 0.75)),
                         float(hp.get("dropouti", 0.2)), float(hp.get("dropoute", 0.1)), genotype=genotype_from_hparams(hp))
        logging.info("Model total parameters: %d", sum(p.numel() for p in model.parameters()))
        self.model = context.wrap_model(model)
        self.base_lr, self.bptt = float(hp.get("learning_rate", 20.0)), int(hp.get("bptt", 35))
        self.optimizer = context.wrap_optimizer(torch.optim.SGD(self.model.parameters(), lr=self.base_lr,
                                                                weight_decay=float(hp.get("weight_decay", 8e-7))))
        self.averaging = AveragedSGD(list(self.model.parameters()))
        clip = float(hp.get("clip_gradients_l2_norm", 0.25))
        self.clip_grads = dp.ClipGradsL2Norm(clip) if clip > 0 else None
        self.alpha, self.beta = float(hp.get("alpha", 0.0)), float(hp.get("beta", 1e-3))
        self.fixed_seq_len = bool(hp.get("fixed_seq_len", False))
        self.hidden: Optional[List[torch.Tensor]] = None
        self._eval_history: List[float] = []
        self._last_loss: Optional[float] = None
        self._last_epoch = -1
        if hp.get("use_cuda_graphs", False):
            if not self.fixed_seq_len:
                raise ValueError("use_cuda_graphs needs fixed_seq_len: true (static shapes)")
            context.experimental.enable_cuda_graphs()

    def build_callbacks(self) -> Dict[str, dp.PyTorchCallback]:
        return {"averaging": _AveragingState(self)}

    def build_training_data_loader(self) -> dp.DataLoader:
        ds = PTBData(self.corpus.train, self.context.get_per_slot_batch_size())
        sampler = BPTTBatchSampler(len(ds), self.bptt, int(self.hp.get("max_seq_length_delta", 20)), fixed=self.fixed_seq_len,
                                   seed=self.context.env.trial_seed)
        return dp.DataLoader(ds, batch_sampler=sampler, collate_fn=split_window)

    def build_validation_data_loader(self) -> dp.DataLoader:
        ds = PTBData(self.corpus.valid, int(self.hp.get("eval_batch_size", 10)))
        return dp.DataLoader(ds, batch_sampler=BPTTBatchSampler(len(ds), self.bptt, 0, valid=True), collate_fn=split_window)

    def _maybe_switch_optimizer(self) -> None:
        nonmono = int(self.hp.get("nonmono", 5))
        if len(self._eval_history) > nonmono + 1 and self._last_loss is not None:
            if self._last_loss > min(self._eval_history[:-(nonmono + 1)]):
                logging.info("Switching to ASGD.")
                self.averaging.set_optim("ASGD")

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
        if epoch_idx != self._last_epoch:
            logging.info("Starting epoch %d", epoch_idx)
            # (the running average's weight changes every step on the host, so it is not offered under graph replay)
            if (epoch_idx > int(self.hp.get("optimizer_switch_epoch", 75)) and self.averaging.optim_name == "SGD"
                    and not self.context.experimental.cuda_graphs):
                self._maybe_switch_optimizer()
            self._last_epoch = epoch_idx
        features, labels = batch
        seq_len, bsz = features.shape
        for g in self.optimizer.param_groups:                      # a longer window is a bigger step
            g["lr"] = self.base_lr * seq_len / self.bptt
        if self.hidden is None or self.hidden[0].shape[1] != bsz:
            self.hidden = self.model.init_hidden(bsz)
        hidden = [h.detach() for h in self.hidden]
        logits, new_hidden, rnn_hs, dropped_rnn_hs = self.model(features, hidden)
        raw_loss = nn.functional.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), labels.reshape(-1))
        loss = raw_loss
        if self.alpha > 0:
            loss = loss + self.alpha * dropped_rnn_hs[-1].float().pow(2).mean()
        if seq_len > 1:
            loss = loss + self.beta * (rnn_hs[-1][1:] - rnn_hs[-1][:-1]).float().pow(2).mean()
        for dst, src in zip(self.hidden, new_hidden):             # in place: the buffer address is stable across graph replays
            dst.copy_(src.detach().to(dst.dtype))
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer, clip_grads=self.clip_grads)
        self.averaging.after_step()
        raw = raw_loss.detach()
        return {"loss": loss.detach(), "raw_loss": raw, "perplexity": torch.exp(raw.clamp(max=11.5))}

    def evaluate_full_dataset(self, data_loader: torch.utils.data.DataLoader) -> Dict[str, Any]:
        self.averaging.swap()
        try:
            hidden = self.model.init_hidden(int(self.hp.get("eval_batch_size", 10)))
            total, rows = torch.zeros((), device=self.context.device), 0
            with torch.no_grad():
                for batch in data_loader:
                    features, targets = self.context.to_device(batch)
                    with self.context.autocast():
                        logits, hidden, _, _ = self.model(features, hidden)
                    total += nn.functional.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), targets.reshape(-1)) * len(features)
                    rows += len(features)
        finally:
            self.averaging.swap()
        mean = float(total) / max(rows, 1)
        perplexity = math.exp(min(mean, 11.5)) if math.isfinite(mean) else 100000.0
        self._last_loss = float(total)
        self._eval_history.append(min([float(total)] + self._eval_history[-1:]))
        return {"loss": float(total), "perplexity": perplexity}
