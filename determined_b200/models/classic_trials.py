"""PyTorch counterparts of the reference examples that exist only for its TensorFlow front-ends. The model families and
training recipes are the reference's; the code is ``PyTorchTrial`` on STOCK ``torch.nn`` modules (these are small models —
the point of the examples is the Trial API, hyperparameter search and the data layer, not the kernels).

* :class:`FashionMNistTrial` — ``examples/tutorials/fashion_mnist_tf_keras/model_def.py`` (Flatten → Dense → Dropout → Dense,
  Adam, sparse categorical cross-entropy, accuracy).
* :class:`IrisTrial` — ``examples/computer_vision/iris_tf_keras/model_def.py`` (two hidden Dense layers sized by
  hyperparameters, RMSprop with learning-rate decay, categorical accuracy as the searcher metric).
* :class:`UNetTrial` — ``examples/computer_vision/unets_tf_keras/model_def.py`` (encoder–decoder with skip connections,
  per-pixel 3-class cross-entropy on 128 x 128 images, pixel accuracy).
* :class:`DataLayerMNistTrial` — ``examples/data_layer/data_layer_mnist_tf_keras/model_def.py``: the data sets come through
  ``context.experimental.cache_train_dataset`` / ``cache_validation_dataset``, so they are built once per version and shared.

No network here: the data sets are synthetic tensors of the real shapes with learnable structure, unless ``data.dir`` names
a directory holding ``train.pt`` / ``validation.pt``.
"""
from __future__ import annotations

import os
from typing import Any, Dict, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import pytorch as dp


def _tensor_pair(context: Any, split: str, make) -> Tuple[torch.Tensor, torch.Tensor]:
    cfg = context.get_data_config() or {}
    d = cfg.get("dir")
    if d and os.path.exists(os.path.join(d, f"{split}.pt")):
        x, y = torch.load(os.path.join(d, f"{split}.pt"))
        return x, y
    return make(int(cfg.get(f"{split}_size", 0)) or None, 1 if split == "train" else 2)


def _loader(context: Any, x: torch.Tensor, y: torch.Tensor, shuffle: bool) -> dp.DataLoader:
    return dp.DataLoader(torch.utils.data.TensorDataset(x, y), batch_size=context.get_per_slot_batch_size(), shuffle=shuffle, drop_last=shuffle)


# ---- Fashion-MNIST tutorial ------------------------------------------------------------------------------------------
def synthetic_images(size: int, seed: int, shape=(1, 28, 28), classes: int = 10) -> Tuple[torch.Tensor, torch.Tensor]:
    g = torch.Generator().manual_seed(seed)
    proto = torch.randn(classes, *shape, generator=torch.Generator().manual_seed(99)) * 0.5
    y = torch.randint(0, classes, (size,), generator=g)
    return torch.randn(size, *shape, generator=g) * 0.5 + proto[y], y


class FashionMNistTrial(dp.PyTorchTrial):
    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        self.model = context.wrap_model(nn.Sequential(nn.Flatten(), nn.Linear(28 * 28, int(hp.get("dense1", 128))), nn.ReLU(),
                                                      nn.Dropout(float(hp.get("dropout", 0.0))), nn.Linear(int(hp.get("dense1", 128)), 10)))
        self.optimizer = context.wrap_optimizer(torch.optim.Adam(self.model.parameters(), lr=float(hp.get("learning_rate", 1e-3))))

    def _data(self, split: str):
        return _tensor_pair(self.context, split, lambda n, seed: synthetic_images(n or (4096 if split == "train" else 1024), seed))

    def build_training_data_loader(self) -> dp.DataLoader:
        return _loader(self.context, *self._data("train"), True)

    def build_validation_data_loader(self) -> dp.DataLoader:
        return _loader(self.context, *self._data("validation"), False)

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        x, y = batch
        logits = self.model(x)
        loss = F.cross_entropy(logits, y)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer)
        return {"loss": loss.detach(), "accuracy": (logits.argmax(-1) == y).float().mean()}

    def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
        x, y = batch
        logits = self.model(x)
        return {"val_loss": F.cross_entropy(logits, y), "val_accuracy": (logits.argmax(-1) == y).float().mean()}


# ---- Iris ------------------------------------------------------------------------------------------------------------
def synthetic_iris(size: int, seed: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Three Gaussian clusters around the species' mean sepal / petal measurements (cm) with the real data's spread."""
    means = torch.tensor([[5.0, 3.4, 1.5, 0.2], [5.9, 2.8, 4.3, 1.3], [6.6, 3.0, 5.6, 2.0]])
    stds = torch.tensor([[0.35, 0.38, 0.17, 0.10], [0.52, 0.31, 0.47, 0.20], [0.64, 0.32, 0.55, 0.27]])
    g = torch.Generator().manual_seed(seed)
    y = torch.randint(0, 3, (size,), generator=g)
    return means[y] + stds[y] * torch.randn(size, 4, generator=g), y


class IrisTrial(dp.PyTorchTrial):
    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        h1, h2 = int(hp.get("layer1_dense_size", 16)), int(hp.get("layer2_dense_size", 16))
        self.model = context.wrap_model(nn.Sequential(nn.Linear(4, h1), nn.ReLU(), nn.Linear(h1, h2), nn.ReLU(), nn.Linear(h2, 3)))
        self.optimizer = context.wrap_optimizer(torch.optim.RMSprop(self.model.parameters(), lr=float(hp.get("learning_rate", 1e-3))))
        # Keras' `decay`: lr / (1 + decay * iterations)
        decay = float(hp.get("learning_rate_decay", 1e-6))
        self.scheduler = context.wrap_lr_scheduler(torch.optim.lr_scheduler.LambdaLR(self.optimizer, lambda it: 1.0 / (1.0 + decay * it)),
                                                   dp.LRScheduler.StepMode.STEP_EVERY_BATCH)

    def _data(self, split: str):
        return _tensor_pair(self.context, split, lambda n, seed: synthetic_iris(n or (120 if split == "train" else 30), seed))

    def build_training_data_loader(self) -> dp.DataLoader:
        return _loader(self.context, *self._data("train"), True)

    def build_validation_data_loader(self) -> dp.DataLoader:
        return _loader(self.context, *self._data("validation"), False)

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        x, y = batch
        logits = self.model(x)
        loss = F.cross_entropy(logits, y)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer)
        return {"loss": loss.detach(), "categorical_accuracy": (logits.argmax(-1) == y).float().mean()}

    def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
        x, y = batch
        logits = self.model(x)
        return {"val_loss": F.cross_entropy(logits, y), "val_categorical_accuracy": (logits.argmax(-1) == y).float().mean()}


# ---- U-Net segmentation ------------------------------------------------------------------------------------------------
class _Down(nn.Sequential):
    def __init__(self, cin: int, cout: int) -> None:
        super().__init__(nn.Conv2d(cin, cout, 3, stride=2, padding=1, bias=False), nn.BatchNorm2d(cout), nn.ReLU(inplace=True),
                         nn.Conv2d(cout, cout, 3, padding=1, bias=False), nn.BatchNorm2d(cout), nn.ReLU(inplace=True))


class _Up(nn.Sequential):          # the pix2pix upsample block of the reference: transposed conv -> norm -> ReLU
    def __init__(self, cin: int, cout: int) -> None:
        super().__init__(nn.ConvTranspose2d(cin, cout, 3, stride=2, padding=1, output_padding=1, bias=False), nn.BatchNorm2d(cout),
                         nn.ReLU(inplace=True))


class UNet(nn.Module):
    """Four strided encoder stages, four transposed-convolution decoder stages, a skip connection at every resolution, and a
    last transposed convolution back to the input size with one channel per class."""

    def __init__(self, classes: int = 3, width: int = 32) -> None:
        super().__init__()
        w = [width, width * 2, width * 4, width * 8, width * 16]
        self.stem = nn.Sequential(nn.Conv2d(3, w[0], 3, stride=2, padding=1, bias=False), nn.BatchNorm2d(w[0]), nn.ReLU(inplace=True))
        self.down = nn.ModuleList([_Down(w[i], w[i + 1]) for i in range(4)])
        self.up = nn.ModuleList([_Up(w[4], w[3]), _Up(2 * w[3], w[2]), _Up(2 * w[2], w[1]), _Up(2 * w[1], w[0])])
        self.last = nn.ConvTranspose2d(2 * w[0], classes, 3, stride=2, padding=1, output_padding=1)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        skips = [self.stem(x)]
        for d in self.down:
            skips.append(d(skips[-1]))
        y = skips.pop()
        for u in self.up:
            y = torch.cat([u(y), skips.pop()], dim=1)
        return self.last(y)


def synthetic_pets(size: int, seed: int, res: int = 128) -> Tuple[torch.Tensor, torch.Tensor]:
    """Images with an elliptical 'pet' (class 0) on a background (class 1) and a border ring (class 2) — the three mask
    classes of the Oxford-IIIT pets task — coloured per class with noise, so the mask is recoverable from the pixels."""
    g = torch.Generator().manual_seed(seed)
    yy, xx = torch.meshgrid(torch.linspace(-1, 1, res), torch.linspace(-1, 1, res), indexing="ij")
    cx, cy = torch.rand(size, 1, 1, generator=g) - 0.5, torch.rand(size, 1, 1, generator=g) - 0.5
    rx, ry = 0.25 + 0.35 * torch.rand(size, 1, 1, generator=g), 0.25 + 0.35 * torch.rand(size, 1, 1, generator=g)
    r = ((xx - cx) / rx) ** 2 + ((yy - cy) / ry) ** 2
    mask = torch.ones(size, res, res, dtype=torch.int64)
    mask[r < 1.0] = 0
    mask[(r >= 1.0) & (r < 1.35)] = 2
    colours = torch.tensor([[0.8, 0.5, 0.3], [0.2, 0.5, 0.3], [0.9, 0.9, 0.9]])
    img = colours[mask].permute(0, 3, 1, 2) + 0.15 * torch.randn(size, 3, res, res, generator=g)
    return img, mask


class UNetTrial(dp.PyTorchTrial):
    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        self.res = int(hp.get("image_size", 128))
        self.model = context.wrap_model(UNet(3, int(hp.get("width", 32))))
        self.optimizer = context.wrap_optimizer(torch.optim.Adam(self.model.parameters(), lr=float(hp.get("learning_rate", 1e-3))))

    def _data(self, split: str):
        return _tensor_pair(self.context, split, lambda n, seed: synthetic_pets(n or (512 if split == "train" else 128), seed, self.res))

    def build_training_data_loader(self) -> dp.DataLoader:
        return _loader(self.context, *self._data("train"), True)

    def build_validation_data_loader(self) -> dp.DataLoader:
        return _loader(self.context, *self._data("validation"), False)

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        x, mask = batch
        logits = self.model(x)
        loss = F.cross_entropy(logits, mask)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer)
        return {"loss": loss.detach(), "accuracy": (logits.argmax(1) == mask).float().mean()}

    def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
        x, mask = batch
        logits = self.model(x)
        pred = logits.argmax(1)
        inter = ((pred == 0) & (mask == 0)).sum().float()
        union = ((pred == 0) | (mask == 0)).sum().float().clamp_min(1.0)
        return {"val_loss": F.cross_entropy(logits, mask), "val_accuracy": (pred == mask).float().mean(), "val_pet_iou": inter / union}


# ---- MNIST through the data layer ----------------------------------------------------------------------------------------
class DataLayerMNistTrial(FashionMNistTrial):
    """The tutorial classifier with both data sets behind the data-layer cache: the first trial (or rank) that asks builds the
    data set under a cluster-wide writer lease, every later one reads the cached copy; ``data.dataset_version`` bumps it."""

    def _cached(self, split: str):
        version = str((self.context.get_data_config() or {}).get("dataset_version", "1"))
        exp = self.context.experimental
        deco = exp.cache_train_dataset("mnist-synthetic", version, shuffle=True) if split == "train" else \
            exp.cache_validation_dataset("mnist-synthetic", version)

        @deco
        def build() -> torch.utils.data.Dataset:
            return torch.utils.data.TensorDataset(*self._data(split))

        return build()

    def build_training_data_loader(self) -> dp.DataLoader:
        return dp.DataLoader(self._cached("train"), batch_size=self.context.get_per_slot_batch_size(), shuffle=True, drop_last=True)

    def build_validation_data_loader(self) -> dp.DataLoader:
        return dp.DataLoader(self._cached("validation"), batch_size=self.context.get_per_slot_batch_size())
