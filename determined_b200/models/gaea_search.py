"""Weight-sharing architecture search with exponentiated-gradient updates (GAEA) over the DARTS cell space.

Counterpart of the reference's ``examples/nas/gaea_pytorch/search`` (``model_search.py``, ``optimizer.py``, ``model_def.py``;
Li et al. 2020, "Geometry-Aware Gradient Algorithms for Neural Architecture Search"): a supernet whose every edge computes
a convex combination of all candidate operations; the mixture weights live ON the probability simplex (no softmax) and are
updated multiplicatively, ``alpha <- alpha * exp(-lr * grad)`` followed by renormalisation, which drives them to sparse
solutions; the shared weights train with SGD. Each batch is split: one half updates the shared weights, the other the
architecture. The discovered cell is read off by keeping, per node, the two strongest incoming edges and their best
non-zero operation — the genotype format ``DARTSCNNTrial`` evaluates (``models/darts.py``).

The candidate operations are the sm_100a modules of ``models/darts.py`` (depthwise + 1x1 GEMM separable / dilated
convolutions, windowed pooling), so on CUDA the supernet runs on the same kernels as the evaluation network.
"""
from __future__ import annotations

from typing import Any, Dict, List, Tuple

import torch
import torch.nn as nn

from .. import pytorch as dp
from ..ops.conv import BatchNorm2d, Conv2d, GlobalAvgPool2d, conv_bn
from ..ops.modules import Linear, cross_entropy
from .darts import PRIMITIVES, FactorizedReduce, Genotype, ReLUConvBN, accuracy_topk, make_op
from .vision_trials import _ClassifierTrial

CANDIDATES = PRIMITIVES[1:]          # "none" is the implicit zero operation: it gets a weight but computes nothing
STEPS = 4
NUM_EDGES = sum(2 + i for i in range(STEPS))   # 14


class MixedOp(nn.Module):
    def __init__(self, c: int, stride: int) -> None:
        super().__init__()
        self.ops = nn.ModuleList(make_op(name, c, stride) for name in CANDIDATES)

    def forward(self, x: torch.Tensor, weights: torch.Tensor) -> torch.Tensor:
        # weights[0] belongs to "none"; the rest follow CANDIDATES
        out = None
        for k, op in enumerate(self.ops):
            y = op(x) * weights[k + 1].to(x.dtype if x.is_cuda else torch.float32)
            out = y if out is None else out + y
        return out


class SearchCell(nn.Module):
    def __init__(self, c_pp: int, c_p: int, c: int, reduction: bool, reduction_prev: bool) -> None:
        super().__init__()
        self.reduction = reduction
        self.pre0 = FactorizedReduce(c_pp, c) if reduction_prev else ReLUConvBN(c_pp, c, 1, 1, 0)
        self.pre1 = ReLUConvBN(c_p, c, 1, 1, 0)
        self.edges = nn.ModuleList()
        for i in range(STEPS):
            for j in range(2 + i):
                self.edges.append(MixedOp(c, 2 if reduction and j < 2 else 1))

    def forward(self, s0: torch.Tensor, s1: torch.Tensor, alphas: torch.Tensor) -> torch.Tensor:
        states = [self.pre0(s0), self.pre1(s1)]
        e = 0
        for i in range(STEPS):
            acc = None
            for j in range(2 + i):
                y = self.edges[e](states[j], alphas[e])
                acc = y if acc is None else acc + y
                e += 1
            states.append(acc)
        return torch.cat(states[-STEPS:], dim=1)


class SearchNetwork(nn.Module):
    """Supernet. ``arch_parameters()`` are two ``[14, 8]`` matrices of simplex rows (normal / reduction cells)."""

    def __init__(self, init_channels: int = 16, num_classes: int = 10, layers: int = 8, stem_multiplier: int = 3) -> None:
        super().__init__()
        self.num_classes = num_classes
        c_curr = stem_multiplier * init_channels
        self.stem_conv = Conv2d(3, c_curr, 3, padding=1, bias=False)
        self.stem_bn = BatchNorm2d(c_curr)
        c_pp, c_p, c_curr = c_curr, c_curr, init_channels
        self.cells = nn.ModuleList()
        reduction_prev = False
        for i in range(layers):
            reduction = layers >= 3 and i in (layers // 3, 2 * layers // 3)
            if reduction:
                c_curr *= 2
            self.cells.append(SearchCell(c_pp, c_p, c_curr, reduction, reduction_prev))
            reduction_prev = reduction
            c_pp, c_p = c_p, STEPS * c_curr
        self.global_pooling = GlobalAvgPool2d()
        self.classifier = Linear(c_p, (num_classes + 7) // 8 * 8)
        k = len(PRIMITIVES)
        self.alphas_normal = nn.Parameter(torch.full((NUM_EDGES, k), 1.0 / k))
        self.alphas_reduce = nn.Parameter(torch.full((NUM_EDGES, k), 1.0 / k))

    def arch_parameters(self) -> List[nn.Parameter]:
        return [self.alphas_normal, self.alphas_reduce]

    def weight_parameters(self) -> List[nn.Parameter]:
        arch = {id(p) for p in self.arch_parameters()}
        return [p for p in self.parameters() if id(p) not in arch]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        s0 = s1 = conv_bn(self.stem_conv, self.stem_bn, x)
        for cell in self.cells:
            s0, s1 = s1, cell(s0, s1, self.alphas_reduce if cell.reduction else self.alphas_normal)
        return self.classifier(self.global_pooling(s1).flatten(1))[:, :self.num_classes]

    def genotype(self) -> Genotype:
        def parse(alphas: torch.Tensor) -> List[Tuple[str, int]]:
            a = alphas.detach().float().cpu()
            gene, start = [], 0
            for i in range(STEPS):
                rows = a[start:start + 2 + i]
                strength = rows[:, 1:].max(dim=1).values                     # best non-zero operation of every incoming edge
                for j in sorted(strength.topk(2).indices.tolist()):
                    gene.append((CANDIDATES[int(rows[j, 1:].argmax())], j))
                start += 2 + i
            return gene

        concat = list(range(2, 2 + STEPS))
        return Genotype(normal=parse(self.alphas_normal), normal_concat=concat, reduce=parse(self.alphas_reduce), reduce_concat=concat)


class ExponentiatedGradient(torch.optim.Optimizer):
    """``p <- p * exp(-lr * grad)``, every row renormalised onto the simplex (reference ``search/optimizer.py``)."""

    def __init__(self, params, lr: float = 0.1, floor: float = 1e-6) -> None:
        super().__init__(params, dict(lr=lr, floor=floor))

    @torch.no_grad()
    def step(self, closure=None):
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                p.mul_(torch.exp(-group["lr"] * p.grad)).clamp_(min=group["floor"])
                p.div_(p.sum(dim=-1, keepdim=True))
        return None


def genotype_to_hparams(g: Genotype) -> Dict[str, Any]:
    """The hyperparameter encoding ``DARTSCNNTrial`` reads (``genotype_from_hparams`` inverts it)."""
    hp: Dict[str, Any] = {}
    for cell, edges in (("normal", g.normal), ("reduce", g.reduce)):
        for n, (op, src) in enumerate(edges):
            i, j = n // 2 + 1, n % 2 + 1
            hp[f"{cell}_node{i}_edge{j}"] = int(src)
            hp[f"{cell}_node{i}_edge{j}_op"] = op
    return hp


class GAEASearchTrial(_ClassifierTrial):
    shape, num_classes = (3, 32, 32), 10
    train_size, val_size = 4096, 1024

    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        self.model = context.wrap_model(SearchNetwork(int(hp.get("init_channels", 16)), self.num_classes, int(hp.get("layers", 8))))
        net = self.model
        self.optimizer = context.wrap_optimizer(torch.optim.SGD(
            net.weight_parameters(), lr=float(hp.get("learning_rate", 0.1)), momentum=float(hp.get("momentum", 0.9)),
            weight_decay=float(hp.get("weight_decay", 3e-4))))
        self.arch_optimizer = context.wrap_optimizer(ExponentiatedGradient(net.arch_parameters(), lr=float(hp.get("arch_learning_rate", 0.1))))
        self.lr_scheduler = context.wrap_lr_scheduler(
            torch.optim.lr_scheduler.CosineAnnealingLR(self.optimizer, int(hp.get("train_epochs", 50)), eta_min=float(hp.get("min_learning_rate", 0.0))),
            dp.LRScheduler.StepMode.STEP_EVERY_EPOCH)
        g = float(hp.get("clip_gradients_l2_norm", 5.0))
        self.clip = dp.ClipGradsL2Norm(g) if g > 0 else None

    def _toggle(self, weights: bool) -> None:
        for p in self.model.weight_parameters():
            p.requires_grad_(weights)
        for p in self.model.arch_parameters():
            p.requires_grad_(not weights)

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        data, labels = batch
        data = self._inputs(data)
        half = data.shape[0] // 2
        # architecture step on the second half of the batch (the reference's held-out split) ...
        self._toggle(weights=False)
        arch_loss = cross_entropy(self.model(data[half:]), labels[half:])
        self.context.backward(arch_loss)
        self.context.step_optimizer(self.arch_optimizer)
        # ... then the shared weights on the first half
        self._toggle(weights=True)
        logits = self.model(data[:half])
        loss = cross_entropy(logits, labels[:half])
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer, clip_grads=self.clip)
        top1, _ = accuracy_topk(logits.detach().float(), labels[:half])
        return {"loss": loss.detach(), "arch_loss": arch_loss.detach(), "top1_accuracy": top1}

    def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
        data, labels = batch
        logits = self.model(self._inputs(data))
        top1, top5 = accuracy_topk(logits.float(), labels)
        ent = -(self.model.alphas_normal * self.model.alphas_normal.clamp_min(1e-12).log()).sum(-1).mean()
        return {"loss": cross_entropy(logits, labels), "top1_accuracy": top1, "top5_accuracy": top5, "arch_entropy": ent}

    def build_callbacks(self) -> Dict[str, dp.PyTorchCallback]:
        trial = self

        class LogGenotype(dp.PyTorchCallback):
            def on_validation_end(self, metrics: Dict[str, Any]) -> None:
                print(f"genotype: {trial.model.genotype()}", flush=True)

        return {"genotype": LogGenotype()}
