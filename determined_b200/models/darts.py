"""DARTS cell networks (CIFAR-10 evaluation network) on the sm_100a modules.

Counterpart of the reference's ``examples/hp_search_benchmarks/darts_cifar10_pytorch`` (``model.py``, ``operations.py``,
``genotypes.py``; the same cell family is used by ``examples/nas/gaea_pytorch``): a stack of normal / reduction cells
whose wiring and operations come from a genotype — in the reference's trial the genotype itself is the hyperparameter
space (``normal_node{i}_edge{j}`` picks the input state, ``..._op`` the operation), which is how ASHA searches
architectures.

Operation set: 3x3 max / average pooling, identity (factorized reduce when strided), separable 3x3 / 5x5 and dilated
3x3 / 5x5 convolutions. Separable and dilated convolutions are a depthwise convolution (``dwconv_ops.cu``) followed by a
1x1 convolution that is a plain tcgen05 GEMM on the NHWC array with the BatchNorm statistics taken from its epilogue.
Keep ``init_channels`` a multiple of 8 so every operation stays on those kernels (the reference's 36 would send the
depthwise layers to torch).
"""
from __future__ import annotations

import collections
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn

from ..ops.conv import Activation, AvgPool2d, BatchNorm2d, Conv2d, DepthwiseConv2d, GlobalAvgPool2d, MaxPool2d, conv_bn
from ..ops.modules import Linear

Genotype = collections.namedtuple("Genotype", "normal normal_concat reduce reduce_concat")

PRIMITIVES = ["none", "max_pool_3x3", "avg_pool_3x3", "skip_connect", "sep_conv_3x3", "sep_conv_5x5", "dil_conv_3x3", "dil_conv_5x5"]

# The second-order DARTS architecture (Liu et al. 2019, "DARTS_V2")
DARTS_V2 = Genotype(
    normal=[("sep_conv_3x3", 0), ("sep_conv_3x3", 1), ("sep_conv_3x3", 0), ("sep_conv_3x3", 1), ("sep_conv_3x3", 1), ("skip_connect", 0),
            ("skip_connect", 0), ("dil_conv_3x3", 2)],
    normal_concat=[2, 3, 4, 5],
    reduce=[("max_pool_3x3", 0), ("max_pool_3x3", 1), ("skip_connect", 2), ("max_pool_3x3", 1), ("max_pool_3x3", 0), ("skip_connect", 2),
            ("skip_connect", 2), ("max_pool_3x3", 1)],
    reduce_concat=[2, 3, 4, 5])


def genotype_from_hparams(hp: Dict[str, Any]) -> Genotype:
    """The reference's encoding (``model_def.py`` ``get_genotype_from_hps``): for node i in 1..4 and edge j in 1..2,
    ``{cell}_node{i}_edge{j}`` is the index of the input state and ``{cell}_node{i}_edge{j}_op`` the primitive. Missing
    keys fall back to DARTS_V2."""
    cells = {}
    for cell, default in (("normal", DARTS_V2.normal), ("reduce", DARTS_V2.reduce)):
        edges: List[Tuple[str, int]] = []
        for i in range(1, 5):
            for j in range(1, 3):
                d_op, d_in = default[2 * (i - 1) + (j - 1)]
                op = str(hp.get(f"{cell}_node{i}_edge{j}_op", d_op))
                src = int(hp.get(f"{cell}_node{i}_edge{j}", d_in))
                if op not in PRIMITIVES or op == "none":
                    raise ValueError(f"unknown operation {op!r}; choose from {PRIMITIVES[1:]}")
                if not 0 <= src < i + 1:
                    raise ValueError(f"{cell} node {i} can only read states 0..{i}, got {src}")
                edges.append((op, src))
        cells[cell] = edges
    return Genotype(normal=cells["normal"], normal_concat=[2, 3, 4, 5], reduce=cells["reduce"], reduce_concat=[2, 3, 4, 5])


# ---- operations -----------------------------------------------------------------------------------------------------
class ReLUConvBN(nn.Module):
    def __init__(self, c_in: int, c_out: int, k: int, stride: int, pad: int) -> None:
        super().__init__()
        self.relu = Activation("relu")
        self.conv = Conv2d(c_in, c_out, k, stride=stride, padding=pad, bias=False)
        self.bn = BatchNorm2d(c_out)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return conv_bn(self.conv, self.bn, self.relu(x))


class DilConv(nn.Module):
    """ReLU → depthwise (dilated) k×k → 1x1 → BN."""

    def __init__(self, c_in: int, c_out: int, k: int, stride: int, pad: int, dilation: int) -> None:
        super().__init__()
        self.relu = Activation("relu")
        self.dw = DepthwiseConv2d(c_in, k, stride=stride, padding=pad, dilation=dilation)
        self.pw = Conv2d(c_in, c_out, 1, bias=False)
        self.bn = BatchNorm2d(c_out)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return conv_bn(self.pw, self.bn, self.dw(self.relu(x)))


class SepConv(nn.Module):
    """Two depthwise-separable convolutions; the first carries the stride."""

    def __init__(self, c_in: int, c_out: int, k: int, stride: int, pad: int) -> None:
        super().__init__()
        self.a = DilConv(c_in, c_in, k, stride, pad, 1)
        self.b = DilConv(c_in, c_out, k, 1, pad, 1)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.b(self.a(x))


class FactorizedReduce(nn.Module):
    """Halve the resolution without losing pixels: two stride-2 1x1 convolutions on the even and the odd grid, concatenated."""

    def __init__(self, c_in: int, c_out: int) -> None:
        super().__init__()
        assert c_out % 2 == 0
        self.relu = Activation("relu")
        self.conv_1 = Conv2d(c_in, c_out // 2, 1, stride=2, bias=False)
        self.conv_2 = Conv2d(c_in, c_out // 2, 1, stride=2, bias=False)
        self.bn = BatchNorm2d(c_out)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.relu(x)
        return self.bn(torch.cat([self.conv_1(x), self.conv_2(x[:, :, 1:, 1:])], dim=1))


class Identity(nn.Module):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return x


class _PoolBN(nn.Module):
    def __init__(self, pool: nn.Module) -> None:
        super().__init__()
        self.pool = pool

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.pool(x)


def make_op(name: str, c: int, stride: int) -> nn.Module:
    if name == "max_pool_3x3":
        return _PoolBN(MaxPool2d(3, stride, 1))
    if name == "avg_pool_3x3":
        return _PoolBN(AvgPool2d(3, stride, 1, count_include_pad=False))
    if name == "skip_connect":
        return Identity() if stride == 1 else FactorizedReduce(c, c)
    if name == "sep_conv_3x3":
        return SepConv(c, c, 3, stride, 1)
    if name == "sep_conv_5x5":
        return SepConv(c, c, 5, stride, 2)
    if name == "dil_conv_3x3":
        return DilConv(c, c, 3, stride, 2, 2)
    if name == "dil_conv_5x5":
        return DilConv(c, c, 5, stride, 4, 2)
    raise ValueError(f"unknown operation {name!r}")


def drop_path(x: torch.Tensor, drop_prob) -> torch.Tensor:
    """Zero whole samples of a branch with probability ``drop_prob`` (scaled so the expectation is unchanged).
    ``drop_prob`` may be a 0-dim device tensor, so a schedule can change it between CUDA-graph replays."""
    keep = 1.0 - drop_prob
    mask = (torch.rand(x.shape[0], 1, 1, 1, device=x.device) < keep).to(x.dtype) / keep
    return x * mask.to(x.dtype)


# ---- cell / network ---------------------------------------------------------------------------------------------------
class Cell(nn.Module):
    def __init__(self, genotype: Genotype, c_pp: int, c_p: int, c: int, reduction: bool, reduction_prev: bool) -> None:
        super().__init__()
        self.reduction = reduction
        self.pre0 = FactorizedReduce(c_pp, c) if reduction_prev else ReLUConvBN(c_pp, c, 1, 1, 0)
        self.pre1 = ReLUConvBN(c_p, c, 1, 1, 0)
        edges, self.concat = (genotype.reduce, list(genotype.reduce_concat)) if reduction else (genotype.normal, list(genotype.normal_concat))
        self.indices = [src for _, src in edges]
        self.ops = nn.ModuleList([make_op(name, c, 2 if reduction and src < 2 else 1) for name, src in edges])
        self.multiplier = len(self.concat)

    def forward(self, s0: torch.Tensor, s1: torch.Tensor, drop_prob: float) -> torch.Tensor:
        states = [self.pre0(s0), self.pre1(s1)]
        for i in range(len(self.ops) // 2):
            hs = []
            for e in (2 * i, 2 * i + 1):
                h = self.ops[e](states[self.indices[e]])
                if self.training and (isinstance(drop_prob, torch.Tensor) or drop_prob > 0.0) and not isinstance(self.ops[e], Identity):
                    h = drop_path(h, drop_prob)
                hs.append(h)
            states.append(hs[0] + hs[1])
        return torch.cat([states[i] for i in self.concat], dim=1)


class AuxiliaryHeadCIFAR(nn.Module):
    """Auxiliary classifier attached two thirds of the way up (input 8x8)."""

    def __init__(self, c: int, num_classes: int) -> None:
        super().__init__()
        self.relu0 = Activation("relu")
        self.pool = AvgPool2d(5, stride=3, padding=0, count_include_pad=False)
        self.conv1 = Conv2d(c, 128, 1, bias=False)
        self.bn1 = BatchNorm2d(128, activation="relu")
        self.conv2 = Conv2d(128, 768, 2, bias=False)
        self.bn2 = BatchNorm2d(768, activation="relu")
        self.classifier = Linear(768, (num_classes + 7) // 8 * 8)
        self.num_classes = num_classes

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = conv_bn(self.conv1, self.bn1, self.pool(self.relu0(x)))
        x = conv_bn(self.conv2, self.bn2, x)
        return self.classifier(x.flatten(1))[:, :self.num_classes]


class DartsNetworkCIFAR(nn.Module):
    """``forward(x)`` → ``(logits, aux_logits or None)``; ``drop_path_prob`` is set by the trial as training proceeds."""

    def __init__(self, init_channels: int = 32, num_classes: int = 10, layers: int = 20, auxiliary: bool = True,
                 genotype: Genotype = DARTS_V2, stem_multiplier: int = 3) -> None:
        super().__init__()
        self.layers, self.auxiliary, self.num_classes = layers, auxiliary, num_classes
        self.drop_path_prob = 0.0
        c_curr = stem_multiplier * init_channels
        self.stem_conv = Conv2d(3, c_curr, 3, padding=1, bias=False)
        self.stem_bn = BatchNorm2d(c_curr)
        c_pp, c_p, c_curr = c_curr, c_curr, init_channels
        self.cells = nn.ModuleList()
        reduction_prev = False
        c_aux = 0
        for i in range(layers):
            reduction = layers >= 3 and i in (layers // 3, 2 * layers // 3)
            if reduction:
                c_curr *= 2
            cell = Cell(genotype, c_pp, c_p, c_curr, reduction, reduction_prev)
            reduction_prev = reduction
            self.cells.append(cell)
            c_pp, c_p = c_p, cell.multiplier * c_curr
            if i == 2 * layers // 3:
                c_aux = c_p
        self.aux_index = 2 * layers // 3
        self.auxiliary_head = AuxiliaryHeadCIFAR(c_aux, num_classes) if auxiliary and c_aux else None
        self.global_pooling = GlobalAvgPool2d()
        self.classifier = Linear(c_p, (num_classes + 7) // 8 * 8)

    def forward(self, x: torch.Tensor) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        aux = None
        s0 = s1 = conv_bn(self.stem_conv, self.stem_bn, x)
        for i, cell in enumerate(self.cells):
            s0, s1 = s1, cell(s0, s1, self.drop_path_prob)
            if i == self.aux_index and self.auxiliary_head is not None and self.training:
                aux = self.auxiliary_head(s1)
        out = self.global_pooling(s1)
        return self.classifier(out.flatten(1))[:, :self.num_classes], aux


def count_parameters_in_mb(model: nn.Module) -> float:
    return sum(p.numel() for n, p in model.named_parameters() if "auxiliary" not in n) / 1e6


def accuracy_topk(logits: torch.Tensor, target: torch.Tensor, topk: Sequence[int] = (1, 5)) -> List[torch.Tensor]:
    k = min(max(topk), logits.shape[1])
    pred = logits.topk(k, dim=1).indices
    hit = pred == target[:, None]
    return [hit[:, :min(t, k)].any(dim=1).float().mean() for t in topk]
