"""Prototypical networks for few-shot classification (Omniglot shape) as a ``PyTorchTrial``.

Counterpart of the reference's ``examples/meta_learning/protonet_omniglot_pytorch/model_def.py`` (Snell et al. 2017: a
4-block conv encoder; a class prototype is the mean embedding of its support examples; queries are classified by a
softmax over negative squared distances to the prototypes; meta-training uses more classes per episode than meta-test;
``global_batch_size`` counts episodes per meta-update). The encoder runs on the sm_100a convolution stack
(Conv2d → fused BatchNorm+ReLU → MaxPool2d, NHWC bf16 on CUDA). Omniglot cannot be downloaded here, so episodes are drawn
from a synthetic alphabet: every class is a fixed random stroke pattern, every example a jittered, noisy rendering of it.

Hyperparameters (the reference's names): ``learning_rate``, ``weight_decay``, ``reduce_every``, ``lr_gamma``,
``global_batch_size``, ``val_batch_size``, ``num_classes_train``, ``num_support_train``, ``num_query_train``,
``num_classes_val``, ``num_support_val``, optional ``num_query_val`` (default 5), ``img_resize_dim``, ``hidden_dim``,
``embedding_dim``. ``data``: ``tasks_per_epoch_train``, ``tasks_per_epoch_val``, ``num_characters`` (synthetic alphabet
size), ``validation_portion``.
"""
from __future__ import annotations

from typing import Any, Dict, Tuple

import torch
import torch.nn as nn

from .. import pytorch as dp
from ..ops.conv import BatchNorm2d, Conv2d, MaxPool2d
from ..ops.modules import cross_entropy


class SyntheticOmniglotTasks(torch.utils.data.Dataset):
    """Item ``i`` is one episode: ``(x_support [C*S,1,D,D], y_support [C*S], x_query [C*Q,1,D,D], y_query [C*Q])`` with
    labels 0..C-1 local to the episode and examples ordered by class, as the reference's loss expects after its argsort."""

    def __init__(self, num_tasks: int, class_ids: torch.Tensor, dim: int, n_classes: int, n_support: int, n_query: int,
                 seed: int = 0, alphabet_seed: int = 1234) -> None:
        if len(class_ids) < n_classes:
            raise ValueError(f"episodes of {n_classes} classes need at least that many characters, got {len(class_ids)}")
        self.num_tasks, self.class_ids, self.dim = num_tasks, class_ids, dim
        self.c, self.s, self.q, self.seed = n_classes, n_support, n_query, seed
        g = torch.Generator().manual_seed(alphabet_seed)
        total = int(class_ids.max().item()) + 1
        # a character = a few random strokes on a coarse grid, upsampled: structured enough that nearby renderings cluster
        coarse = (torch.rand(total, 1, 7, 7, generator=g) < 0.3).float()
        self.glyphs = torch.nn.functional.interpolate(coarse, size=(dim, dim), mode="nearest")

    def __len__(self) -> int:
        return self.num_tasks

    def _render(self, cls: torch.Tensor, g: torch.Generator) -> torch.Tensor:
        x = self.glyphs[cls]                                           # [N, 1, D, D]
        sh = torch.randint(-2, 3, (x.shape[0], 2), generator=g)
        x = torch.stack([torch.roll(x[i], (int(sh[i, 0]), int(sh[i, 1])), dims=(1, 2)) for i in range(x.shape[0])])
        flip = (torch.rand(x.shape, generator=g) < 0.04).float()
        return (x * (1 - flip) + (1 - x) * flip) * 2.0 - 1.0

    def __getitem__(self, i: int) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
        g = torch.Generator().manual_seed(self.seed * 1_000_003 + i)
        pick = self.class_ids[torch.randperm(len(self.class_ids), generator=g)[:self.c]]
        ys = torch.arange(self.c).repeat_interleave(self.s)
        yq = torch.arange(self.c).repeat_interleave(self.q)
        return self._render(pick[ys], g), ys, self._render(pick[yq], g), yq


def squared_distances(x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """[N, D], [M, D] → [N, M] squared Euclidean distances (the expansion form: one GEMM instead of an N×M×D tensor)."""
    x, y = x.float(), y.float()
    return (x * x).sum(1, keepdim=True) - 2.0 * x @ y.t() + (y * y).sum(1)[None, :]


class ProtoNetEncoder(nn.Module):
    def __init__(self, x_dim: int = 1, hid_dim: int = 64, z_dim: int = 64) -> None:
        super().__init__()
        dims = [x_dim, hid_dim, hid_dim, hid_dim, z_dim]
        blocks = []
        for i in range(4):
            blocks += [Conv2d(dims[i], dims[i + 1], 3, padding=1), BatchNorm2d(dims[i + 1], activation="relu"), MaxPool2d(2)]
        self.blocks = nn.Sequential(*blocks)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.blocks(x).flatten(1)


def prototypical_loss(embed: nn.Module, xs: torch.Tensor, ys: torch.Tensor, xq: torch.Tensor, yq: torch.Tensor,
                      n_classes: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """One episode: support and query go through the encoder in ONE forward pass (batch statistics shared, as in the
    reference), prototypes are per-class means, the loss is cross-entropy over -distance logits."""
    n_support = ys.shape[0] // n_classes
    z = embed(torch.cat([xs, xq], 0))
    z = z.float()
    protos = z[: n_classes * n_support].view(n_classes, n_support, -1).mean(1)
    logits = -squared_distances(z[n_classes * n_support:], protos)
    loss = cross_entropy(logits, yq)
    acc = (logits.argmax(1) == yq).float().mean()
    return loss, acc


class OmniglotProtoNetTrial(dp.PyTorchTrial):
    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        self.data_config = context.get_data_config() or {}
        self.n_cls = {"train": int(hp.get("num_classes_train", 60)), "val": int(hp.get("num_classes_val", 20))}
        self.n_sup = {"train": int(hp.get("num_support_train", 1)), "val": int(hp.get("num_support_val", 1))}
        self.n_qry = {"train": int(hp.get("num_query_train", 5)), "val": int(hp.get("num_query_val", 5))}
        self.dim = int(hp.get("img_resize_dim", 28))
        n_chars = int(self.data_config.get("num_characters", 1623))    # Omniglot has 1623 characters
        perm = torch.randperm(n_chars, generator=torch.Generator().manual_seed(17))
        n_val = max(self.n_cls["val"], int(n_chars * float(self.data_config.get("validation_portion", 0.25))))
        self.val_classes, self.train_classes = perm[:n_val], perm[n_val:]
        self.model = context.wrap_model(ProtoNetEncoder(1, int(hp.get("hidden_dim", 64)), int(hp.get("embedding_dim", 64))))
        self.optimizer = context.wrap_optimizer(torch.optim.Adam(self.model.parameters(), lr=float(hp.get("learning_rate", 1e-3)),
                                                                 weight_decay=float(hp.get("weight_decay", 0.0))))
        self.lr_scheduler = context.wrap_lr_scheduler(
            torch.optim.lr_scheduler.StepLR(self.optimizer, int(hp.get("reduce_every", 200)), gamma=float(hp.get("lr_gamma", 0.5))),
            dp.LRScheduler.StepMode.STEP_EVERY_EPOCH)

    def _tasks(self, split: str, seed: int) -> SyntheticOmniglotTasks:
        n = int(self.data_config.get("tasks_per_epoch_train" if split == "train" else "tasks_per_epoch_val", 100 if split == "train" else 200))
        classes = self.train_classes if split == "train" else self.val_classes
        return SyntheticOmniglotTasks(n, classes, self.dim, self.n_cls[split], self.n_sup[split], self.n_qry[split], seed)

    def build_training_data_loader(self) -> dp.DataLoader:
        return dp.DataLoader(self._tasks("train", 1), batch_size=self.context.get_per_slot_batch_size(), shuffle=True, drop_last=True,
                             pin_memory=torch.cuda.is_available())

    def build_validation_data_loader(self) -> dp.DataLoader:
        bs = int(self.context.get_hparams().get("val_batch_size", 2))
        return dp.DataLoader(self._tasks("val", 2), batch_size=bs, drop_last=True, pin_memory=torch.cuda.is_available())

    def _episodes(self, batch: Any, split: str) -> Tuple[torch.Tensor, torch.Tensor]:
        xs, ys, xq, yq = batch
        loss = acc = 0.0
        for t in range(xs.shape[0]):        # episodes are independent forward passes: BatchNorm statistics are per episode
            l, a = prototypical_loss(self.model, xs[t], ys[t], xq[t], yq[t], self.n_cls[split])
            loss, acc = loss + l, acc + a
        return loss / xs.shape[0], acc / xs.shape[0]

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        loss, acc = self._episodes(batch, "train")
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer)
        return {"loss": loss.detach(), "acc": acc.detach()}

    def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
        loss, acc = self._episodes(batch, "val")
        return {"loss": loss, "acc": acc}
