"""DARTS CIFAR-10 evaluation network as a ``PyTorchTrial`` whose ARCHITECTURE is a hyperparameter.

Counterpart of the reference's ``examples/hp_search_benchmarks/darts_cifar10_pytorch/model_def.py`` (``DARTSCNNTrial``):
SGD with momentum and weight decay, cosine annealing stepped every epoch, gradient-norm clipping, an auxiliary classifier
weighted by ``auxiliary_weight``, drop-path probability ramped linearly over training, top-1 / top-5 accuracy as
validation metrics; ``adaptive.yaml`` searches over the genotype keys with ASHA (the "NAS as hyperparameter search"
benchmark of the ASHA paper). CIFAR-10 cannot be downloaded here: batches are synthetic images of the CIFAR shape unless
``data.dir`` provides tensors (see ``vision_trials``).

Hyperparameters (the reference's names): ``learning_rate``, ``momentum``, ``weight_decay``, ``train_epochs``,
``global_batch_size``, ``init_channels``, ``layers``, ``auxiliary``, ``auxiliary_weight``, ``cutout``, ``cutout_length``,
``drop_path_prob``, ``clip_gradients_l2_norm``, ``use_cuda_graphs``, and the genotype keys ``{normal,reduce}_node{1..4}_edge{1,2}[_op]``.
"""
from __future__ import annotations

from typing import Any, Dict

import torch

from .. import pytorch as dp
from ..ops.modules import cross_entropy
from .darts import DartsNetworkCIFAR, accuracy_topk, genotype_from_hparams
from .vision_trials import _ClassifierTrial


def cutout_(images: torch.Tensor, length: int, generator: torch.Generator = None) -> torch.Tensor:
    """Zero one random ``length`` × ``length`` square per image, in place (DeVries & Taylor 2017), on whatever device the
    batch lives on."""
    n, _, h, w = images.shape
    cy = torch.randint(0, h, (n,), device=images.device, generator=generator)
    cx = torch.randint(0, w, (n,), device=images.device, generator=generator)
    ys = torch.arange(h, device=images.device)[None, :, None]
    xs = torch.arange(w, device=images.device)[None, None, :]
    hole = ((ys - cy[:, None, None]).abs() <= length // 2) & ((xs - cx[:, None, None]).abs() <= length // 2)
    return images.masked_fill_(hole[:, None, :, :], 0.0)


class DARTSCNNTrial(_ClassifierTrial):
    shape, num_classes = (3, 32, 32), 10
    train_size, val_size = 4096, 1024

    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        self.hp = hp
        self.model = context.wrap_model(DartsNetworkCIFAR(
            int(hp.get("init_channels", 32)), self.num_classes, int(hp.get("layers", 20)), bool(hp.get("auxiliary", True)),
            genotype_from_hparams(hp)))
        self.optimizer = context.wrap_optimizer(torch.optim.SGD(
            self.model.parameters(), lr=float(hp.get("learning_rate", 0.025)), momentum=float(hp.get("momentum", 0.9)),
            weight_decay=float(hp.get("weight_decay", 3e-4))))
        self.train_epochs = int(hp.get("train_epochs", 300))
        self.lr_scheduler = context.wrap_lr_scheduler(
            torch.optim.lr_scheduler.CosineAnnealingLR(self.optimizer, self.train_epochs), dp.LRScheduler.StepMode.STEP_EVERY_EPOCH)
        g = float(hp.get("clip_gradients_l2_norm", 5.0))
        self.clip = dp.ClipGradsL2Norm(g) if g > 0 else None
        self.aux_weight = float(hp.get("auxiliary_weight", 0.4))
        self.drop_path_prob = float(hp.get("drop_path_prob", 0.2))
        self.graphs = bool(hp.get("use_cuda_graphs", False))
        if self.graphs:
            context.experimental.enable_cuda_graphs()

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        data, labels = batch
        data = self._inputs(data)
        if self.hp.get("cutout", True):
            data = cutout_(data.clone(), int(self.hp.get("cutout_length", 16)))
        net = self.model
        if self.graphs and self.drop_path_prob > 0:    # the ramp must survive graph replays: read the epoch from device memory
            epoch = self.context.experimental.step_info()[0]
            net.drop_path_prob = self.drop_path_prob * torch.clamp(epoch / max(self.train_epochs, 1), max=1.0)
        else:
            net.drop_path_prob = self.drop_path_prob * min(epoch_idx, self.train_epochs) / max(self.train_epochs, 1)
        logits, aux = net(data)
        loss = cross_entropy(logits, labels)
        if aux is not None:
            loss = loss + self.aux_weight * cross_entropy(aux, labels)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer, clip_grads=self.clip)
        top1, top5 = accuracy_topk(logits.detach().float(), labels)
        return {"loss": loss.detach(), "top1_accuracy": top1, "top5_accuracy": top5}

    def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
        data, labels = batch
        logits, _ = self.model(self._inputs(data))
        top1, top5 = accuracy_topk(logits.float(), labels)
        return {"loss": cross_entropy(logits, labels), "top1_accuracy": top1, "top5_accuracy": top5}
