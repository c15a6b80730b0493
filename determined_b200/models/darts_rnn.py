"""DARTS recurrent cell language model (Penn Treebank search space of "Random search and reproducibility for NAS").

Counterpart of the reference's ``examples/hp_search_benchmarks/darts_penntreebank_pytorch`` (``randomNAS_files/model.py:15-182``
cell + model, ``utils.py:52-101`` dropouts, ``optimizer.py:6-75`` SGD ↔ ASGD switch). The math is the published cell: a
highway-style initial state ``s0 = h + σ(c0) ⊙ (f(h0) - h)`` with ``[c0, h0] = [x, h] W0``, then eight nodes
``s_i = s_p + σ(c) ⊙ (f_i(h) - s_p)`` with ``[c, h] = s_p W_i`` for the node's predecessor ``p`` and activation ``f_i``, and
the mean of the eight node states as the output; variational ("locked") dropout masks on the input, hidden and output
activations, row dropout on the embedding matrix, and a decoder that shares the embedding matrix.

What is different is the schedule, which is what decides the speed of a strictly sequential recurrence on a large GPU:

* the input half of ``W0`` does not depend on the recurrence (the input dropout mask is the same at every time step), so
  ``x W0[:ninp]`` is ONE ``[T·B, ninp] × [ninp, 2·nhid]`` GEMM per sequence instead of ``T`` small ones;
* nodes that share a predecessor read the same state, so their weight matrices are concatenated once per sequence and
  each group is one GEMM per time step — the default genotype needs 5 recurrent GEMMs per step instead of 9;
* on a GPU the gate of all nodes of such a group — sigmoid, activation, highway mix, the dropout-masked copy that feeds
  the next GEMM — is ONE kernel forward and one backward (``csrc/kernels/rnn_ops.cu``) instead of five elementwise
  launches per node forward and a dozen backward: at batch 64 × width 850 every one of those moves ~100 KB, so the
  recurrence is bound by launch and dependency latency;
* the trial can fix the sequence length (``fixed_seq_len``) so that the whole training step has static shapes and replays
  as a CUDA graph, which removes the launch overhead that otherwise dominates ``T × 5`` small GEMMs.

The hidden width of the published configuration (850) is not a multiple of 8, which the TMA descriptors of the in-tree GEMM
need, so the GEMMs here are torch's (cuBLAS, bf16 under autocast); ``DET_RNN_FUSED_GATE=0`` selects the all-torch path.
"""
from __future__ import annotations

import os
from collections import namedtuple
from typing import Dict, List, Optional, Sequence, Tuple

import torch
from torch import nn

Genotype = namedtuple("Genotype", "recurrent concat")

INITRANGE = 0.04
ACTIVATIONS = {"tanh": torch.tanh, "relu": torch.relu, "sigmoid": torch.sigmoid, "identity": lambda x: x}

# the architecture found by DARTS (second order) on Penn Treebank
DARTS_V2 = Genotype(recurrent=[("sigmoid", 0), ("relu", 1), ("relu", 1), ("identity", 1), ("tanh", 2), ("sigmoid", 5),
                               ("tanh", 3), ("relu", 5)], concat=range(1, 9))


def genotype_from_hparams(hp: Dict[str, object], steps: int = 8) -> Genotype:
    """``node{i}_edge`` / ``node{i}_op`` hyperparameters → genotype (reference ``model_def.py:108-118``); a node may only
    read a state that exists before it."""
    cell = []
    for node in range(steps):
        edge, op = int(hp[f"node{node + 1}_edge"]), str(hp[f"node{node + 1}_op"])
        if not 0 <= edge <= node:
            raise ValueError(f"node{node + 1}_edge must be in [0, {node}]")
        if op not in ACTIVATIONS:
            raise ValueError(f"node{node + 1}_op must be one of {sorted(ACTIVATIONS)}")
        cell.append((op, edge))
    return Genotype(recurrent=cell, concat=range(1, steps + 1))


def locked_mask(like: torch.Tensor, batch: int, width: int, drop: float) -> torch.Tensor:
    """Variational dropout mask ``[batch, width]`` (shared by every time step), scaled by ``1 / keep``."""
    keep = 1.0 - drop
    return torch.bernoulli(torch.full((batch, width), keep, device=like.device, dtype=torch.float32)).to(like.dtype) / keep


class _HighwayGateFn(torch.autograd.Function):
    """``(ch [B, k*2*nh], sp [B, nh], mask) → k node states (+ k masked copies)`` through the fused gate kernels."""

    @staticmethod
    def forward(ctx, ch, sp, mask, acts):
        from ..ops import functional as F

        states, masked = F.highway_gate_fwd(ch, sp, mask, acts)
        ctx.save_for_backward(ch, sp, mask)
        ctx.acts = acts
        ctx.set_materialize_grads(False)
        return tuple(states + masked)

    @staticmethod
    def backward(ctx, *grads):
        from ..ops import functional as F

        ch, sp, mask = ctx.saved_tensors
        k = len(ctx.acts)
        grads = [None if g is None else g.contiguous() for g in grads]
        dch, dsp = F.highway_gate_bwd(ch, sp, mask, grads[:k], grads[k:] if mask is not None else [None] * k, ctx.acts)
        return dch, dsp, None, None


def fused_gate_enabled(x: torch.Tensor, nhid: int) -> bool:
    return x.is_cuda and x.dtype == torch.bfloat16 and nhid % 2 == 0 and os.environ.get("DET_RNN_FUSED_GATE", "1") != "0"


class DARTSCell(nn.Module):
    def __init__(self, ninp: int, nhid: int, dropouth: float, dropoutx: float, genotype: Genotype, init_op: str = "tanh") -> None:
        super().__init__()
        self.ninp, self.nhid, self.dropouth, self.dropoutx = ninp, nhid, dropouth, dropoutx
        self.genotype, self.init_op = genotype, init_op
        self._W0 = nn.Parameter(torch.empty(ninp + nhid, 2 * nhid).uniform_(-INITRANGE, INITRANGE))
        self._Ws = nn.ParameterList(nn.Parameter(torch.empty(nhid, 2 * nhid).uniform_(-INITRANGE, INITRANGE))
                                    for _ in genotype.recurrent)
        groups: Dict[int, List[int]] = {}
        for i, (_, pred) in enumerate(genotype.recurrent):
            groups.setdefault(int(pred), []).append(i)
        # ascending predecessor order is a valid schedule: a node's predecessor always has a smaller index than the node
        self.groups: List[Tuple[int, List[int]]] = sorted(groups.items())

    def forward(self, inputs: torch.Tensor, hidden: torch.Tensor, masks: Optional[Tuple[torch.Tensor, torch.Tensor]] = None,
                fused: Optional[bool] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        """``inputs`` [T, B, ninp], ``hidden`` [1, B, nhid] → (all hidden states [T, B, nhid], last one [1, B, nhid]).
        ``masks = (x_mask, h_mask)`` overrides the sampled dropout masks and ``fused`` the choice of gate path (tests)."""
        T, B, _ = inputs.shape
        x_mask = h_mask = None
        if masks is not None:
            x_mask, h_mask = masks
        elif self.training:
            x_mask = locked_mask(inputs, B, self.ninp, self.dropoutx)
            h_mask = locked_mask(inputs, B, self.nhid, self.dropouth)
        nh = self.nhid
        w0 = self._W0.to(inputs.dtype)
        xin = inputs if x_mask is None else inputs * x_mask
        xproj = (xin.reshape(T * B, self.ninp) @ w0[:self.ninp]).view(T, B, 2 * nh)      # hoisted out of the recurrence
        w0h = w0[self.ninp:]
        wcat = [torch.cat([self._Ws[i] for i in members], dim=1).to(inputs.dtype) for _, members in self.groups]
        acts = [ACTIVATIONS[name] for name, _ in self.genotype.recurrent]
        f0 = ACTIVATIONS[self.init_op]
        n_states = len(self.genotype.recurrent) + 1
        concat = list(self.genotype.concat)
        h = hidden[0].to(inputs.dtype).contiguous()
        out = []
        if fused_gate_enabled(inputs, nh) if fused is None else fused:
            from ..ops import _native
            from ..ops.functional import GATE_ACTS

            _native.load()
            if h_mask is not None:
                h_mask = h_mask.to(inputs.dtype).contiguous()
            init_act = (GATE_ACTS[self.init_op],)
            group_acts = [tuple(GATE_ACTS[self.genotype.recurrent[i][0]] for i in members) for _, members in self.groups]
            for t in range(T):
                hm = h if h_mask is None else h * h_mask
                res = _HighwayGateFn.apply(torch.addmm(xproj[t], hm, w0h), h, h_mask, init_act)
                states = [res[0]] + [None] * (n_states - 1)
                feeds = [res[-1]] + [None] * (n_states - 1)        # what the next GEMM reads: the masked copy, if any
                for (pred, members), w, ga in zip(self.groups, wcat, group_acts):
                    res = _HighwayGateFn.apply(feeds[pred] @ w, states[pred], h_mask, ga)
                    k = len(members)
                    for j, i in enumerate(members):
                        states[i + 1], feeds[i + 1] = res[j], res[-k + j]
                h = torch.stack([states[i] for i in concat], dim=-1).mean(-1)
                out.append(h)
            hiddens = torch.stack(out)
            return hiddens, hiddens[-1].unsqueeze(0)
        for t in range(T):
            hm = h if h_mask is None else h * h_mask
            c0, h0 = (xproj[t] + hm @ w0h).split(nh, dim=-1)
            states: List[Optional[torch.Tensor]] = [None] * n_states
            states[0] = h + torch.sigmoid(c0) * (f0(h0) - h)
            for (pred, members), w in zip(self.groups, wcat):
                sp = states[pred]
                ch = (sp if h_mask is None else sp * h_mask) @ w
                for j, i in enumerate(members):
                    c, hh = ch[:, 2 * nh * j:2 * nh * (j + 1)].split(nh, dim=-1)
                    states[i + 1] = sp + torch.sigmoid(c) * (acts[i](hh) - sp)
            h = torch.stack([states[i] for i in concat], dim=-1).mean(-1)
            out.append(h)
        hiddens = torch.stack(out)
        return hiddens, hiddens[-1].unsqueeze(0)


class RNNModel(nn.Module):
    """Embedding → DARTS cell → tied decoder. ``forward(tokens [T, B], hidden)`` returns ``(logits [T, B, ntoken], hidden,
    raw_outputs, dropped_outputs)``; the last two feed the activation regularisers."""

    def __init__(self, ntoken: int, ninp: int, nhid: int, nhidlast: int, dropout: float = 0.5, dropouth: float = 0.5,
                 dropoutx: float = 0.5, dropouti: float = 0.5, dropoute: float = 0.1, genotype: Genotype = DARTS_V2,
                 init_op: str = "tanh") -> None:
        super().__init__()
        if not ninp == nhid == nhidlast:
            raise ValueError("the tied decoder needs emsize == nhid == nhidlast")
        self.encoder = nn.Embedding(ntoken, ninp)
        self.rnns = nn.ModuleList([DARTSCell(ninp, nhid, dropouth, dropoutx, genotype, init_op)])
        self.decoder = nn.Linear(ninp, ntoken)
        self.decoder.weight = self.encoder.weight
        self.encoder.weight.data.uniform_(-INITRANGE, INITRANGE)
        self.decoder.bias.data.zero_()
        self.ntoken, self.ninp, self.nhid = ntoken, ninp, nhid
        self.dropout, self.dropouti, self.dropoute = dropout, dropouti, dropoute

    def init_hidden(self, bsz: int) -> List[torch.Tensor]:
        w = self.encoder.weight
        return [w.new_zeros(1, bsz, self.nhid) for _ in self.rnns]

    def _lockdrop(self, x: torch.Tensor, drop: float) -> torch.Tensor:
        if not self.training or not drop:
            return x
        return x * locked_mask(x, x.shape[1], x.shape[2], drop)[None]

    def forward(self, tokens: torch.Tensor, hidden: Sequence[torch.Tensor]):
        w = self.encoder.weight
        if self.training and self.dropoute:
            keep = 1.0 - self.dropoute
            w = w * (torch.bernoulli(torch.full((w.shape[0], 1), keep, device=w.device)) / keep)     # whole words dropped
        emb = self._lockdrop(nn.functional.embedding(tokens, w), self.dropouti)
        if emb.is_cuda:
            emb = emb.to(torch.bfloat16)      # bf16 activations and GEMM operands on the GPU; weights stay fp32 masters
        raw = emb
        new_hidden, raw_outputs = [], []
        for cell, h in zip(self.rnns, hidden):
            raw, new_h = cell(raw, h)
            new_hidden.append(new_h)
            raw_outputs.append(raw)
        output = self._lockdrop(raw, self.dropout)
        logits = nn.functional.linear(output, self.decoder.weight.to(output.dtype), self.decoder.bias.to(output.dtype))
        return logits, new_hidden, raw_outputs, [output]


class AveragedSGD:
    """The reference's ``HybridSGD`` (SGD that can be switched to ASGD with ``lambd = 0, t0 = 0``) restated as what it
    computes: with those constants ASGD takes the SAME step as SGD and additionally keeps the running mean ``ax`` of the
    iterates, which replaces the weights during evaluation. Keeping the update itself in a plain ``torch.optim.SGD``
    leaves it on the fused flat-bucket optimizer path; this object only owns the average."""

    def __init__(self, params: Sequence[torch.Tensor]) -> None:
        self.params = list(params)
        self.optim_name = "SGD"
        self.ax: List[torch.Tensor] = []
        self.steps = 0

    def set_optim(self, name: str) -> None:
        if name not in ("SGD", "ASGD"):
            raise ValueError(name)
        if name == "ASGD" and self.optim_name != "ASGD":
            self.ax = [p.detach().clone() for p in self.params]
            self.steps = 0
        self.optim_name = name

    @torch.no_grad()
    def after_step(self) -> None:
        if self.optim_name != "ASGD":
            return
        self.steps += 1
        mu = 1.0 / max(1, self.steps)
        torch._foreach_lerp_(self.ax, [p.detach() for p in self.params], mu)

    @torch.no_grad()
    def swap(self) -> None:
        """Exchange weights and their average (call before and after evaluation)."""
        if self.optim_name != "ASGD":
            return
        for p, a in zip(self.params, self.ax):
            tmp = p.detach().clone()
            p.detach().copy_(a)
            a.copy_(tmp)

    def state_dict(self) -> Dict[str, object]:
        return {"optim_name": self.optim_name, "steps": self.steps, "ax": [a.cpu() for a in self.ax]}

    def load_state_dict(self, sd: Dict[str, object]) -> None:
        self.optim_name, self.steps = str(sd["optim_name"]), int(sd["steps"])          # type: ignore[arg-type]
        self.ax = [a.to(p.device) for a, p in zip(sd["ax"], self.params)]              # type: ignore[arg-type]
