"""``PyTorchTrial`` definitions for the vision / GAN side of the example zoo.

Each trial mirrors the training recipe of its reference example; the data sets are synthetic tensors of the real
shapes (no network in this environment) unless ``data.dir`` in the experiment config points at prepared ``.pt`` files.

* :class:`MNistTrial`  — reference ``examples/tutorials/mnist_pytorch/model_def.py:27-94`` (Adadelta, NLL loss).
* :class:`CIFARTrial`  — reference ``examples/computer_vision/cifar10_pytorch/model_def.py:38-137`` (RMSprop with decay,
  ``validation_error`` as searcher metric).
* :class:`ResNet50Trial` — BASELINE config #2, authored new (SGD momentum 0.9, step LR, synthetic ImageNet batches).
* :class:`GANTrial` — reference ``examples/gan/gan_mnist_pytorch/model_def.py:74-184`` (two models, two Adam
  optimizers, ``requires_grad_`` toggling, LambdaLR stepped every epoch).
* :class:`DCGANTrial` — the same multi-model / multi-optimizer pattern on the convolutional DCGAN pair (BASELINE config #4).
"""
from __future__ import annotations

import os
from typing import Any, Dict, Tuple

import torch
import torch.nn as nn

from .. import pytorch as dp
from ..ops.modules import cross_entropy
from .resnet import resnet18, resnet50
from .vision import CifarCNN, DCGANDiscriminator, DCGANGenerator, MLPDiscriminator, MLPGenerator, MnistCNN, dcgan_init


class SyntheticImageDataset(torch.utils.data.Dataset):
    """``size`` random images of ``shape`` with labels in ``[0, num_classes)``; class-dependent mean so that a network
    can actually learn something from it (losses go down, validation error moves)."""

    det_batched = True   # __getitems__ returns a collated batch (one gather instead of per-sample indexing + stack)

    def __init__(self, size: int, shape: Tuple[int, int, int], num_classes: int, seed: int = 0, uint8_nhwc: bool = False) -> None:
        g = torch.Generator().manual_seed(seed)
        self.labels = torch.randint(0, num_classes, (size,), generator=g, dtype=torch.int64)
        if uint8_nhwc:
            # decoded-image layout: bytes, height x width x channels (what a JPEG decoder hands over); low-resolution
            # class prototypes keep construction cheap at ImageNet sizes
            C, H, W = shape
            proto = torch.randint(64, 192, (num_classes, 8, 8, C), generator=torch.Generator().manual_seed(1234), dtype=torch.int16)
            noise = torch.randint(-48, 48, (size, H, W, C), generator=g, dtype=torch.int16)
            up = proto[self.labels].repeat_interleave((H + 7) // 8, 1).repeat_interleave((W + 7) // 8, 2)[:, :H, :W]
            self.images = (up + noise).clamp_(0, 255).to(torch.uint8)
        else:
            proto = torch.randn(num_classes, *shape, generator=torch.Generator().manual_seed(1234)) * 0.5
            self.images = torch.randn(size, *shape, generator=g) * 0.5 + proto[self.labels]

    def __len__(self) -> int:
        return self.images.shape[0]

    def __getitem__(self, i: int):
        return self.images[i], self.labels[i]

    def __getitems__(self, idxs):
        idx = torch.as_tensor(idxs, dtype=torch.int64)
        return self.images.index_select(0, idx), self.labels.index_select(0, idx)


IMAGENET_MEAN, IMAGENET_STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)


def normalize_uint8_images(data: torch.Tensor, mean=IMAGENET_MEAN, std=IMAGENET_STD) -> torch.Tensor:
    """uint8 NHWC bytes → normalised logical-NCHW activations: one fused kernel on CUDA (bf16 NHWC, no transpose needed
    because the convolutions work in NHWC), plain torch on CPU."""
    if data.is_cuda:
        from ..ops import functional as F
        from ..ops.conv import from_nhwc

        return from_nhwc(F.u8_nhwc_to_bf16(data.contiguous(), mean, std))
    m = torch.tensor(mean).view(1, -1, 1, 1)
    s_ = torch.tensor(std).view(1, -1, 1, 1)
    return (data.permute(0, 3, 1, 2).float() / 255.0 - m) / s_


def _dataset(context, split: str, default_size: int, shape, num_classes: int, uint8_nhwc: bool = False) -> torch.utils.data.Dataset:
    data_cfg = context.get_data_config() or {}
    d = data_cfg.get("dir")
    if d and os.path.exists(os.path.join(d, f"{split}.pt")):
        images, labels = torch.load(os.path.join(d, f"{split}.pt"))
        return torch.utils.data.TensorDataset(images.float(), labels.long())
    size = int(data_cfg.get(f"{split}_size", default_size))
    return SyntheticImageDataset(size, shape, num_classes, seed=1 if split == "train" else 2, uint8_nhwc=uint8_nhwc)


def _loader(context, ds, shuffle: bool, prefetch: int = 0) -> dp.DataLoader:
    return dp.DataLoader(ds, batch_size=context.get_per_slot_batch_size(), shuffle=shuffle, drop_last=True,
                         pin_memory=torch.cuda.is_available(), prefetch_batches=prefetch)


class _ClassifierTrial(dp.PyTorchTrial):
    shape: Tuple[int, int, int] = (1, 28, 28)
    num_classes = 10
    train_size, val_size = 4096, 1024
    uint8_input = False   # data set serves uint8 NHWC bytes; normalisation runs on the device
    prefetch = 0

    def build_training_data_loader(self) -> dp.DataLoader:
        ds = _dataset(self.context, "train", self.train_size, self.shape, self.num_classes, self.uint8_input)
        return _loader(self.context, ds, True, self.prefetch)

    def build_validation_data_loader(self) -> dp.DataLoader:
        ds = _dataset(self.context, "validation", self.val_size, self.shape, self.num_classes, self.uint8_input)
        return _loader(self.context, ds, False)

    def _inputs(self, data: torch.Tensor) -> torch.Tensor:
        return normalize_uint8_images(data) if data.dtype == torch.uint8 else data

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        data, labels = batch
        loss = cross_entropy(self.model(self._inputs(data)), labels)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer)
        return {"loss": loss.detach()}

    def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
        data, labels = batch
        logits = self.model(self._inputs(data))
        loss = cross_entropy(logits, labels)
        acc = (logits.argmax(dim=-1) == labels).float().mean()
        return {"validation_loss": loss, "accuracy": acc, "validation_error": 1.0 - acc}


class MNistTrial(_ClassifierTrial):
    shape, num_classes = (1, 28, 28), 10

    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        self.model = context.wrap_model(MnistCNN(int(hp.get("n_filters1", 32)), int(hp.get("n_filters2", 64)),
                                                 float(hp.get("dropout1", 0.25)), float(hp.get("dropout2", 0.5))))
        self.optimizer = context.wrap_optimizer(torch.optim.Adadelta(self.model.parameters(), lr=float(hp.get("learning_rate", 1.0))))
        if hp.get("overlap_grad_allreduce", False):
            context.experimental.enable_grad_overlap(chunk_mb=int(hp.get("overlap_chunk_mb", 1)))
        if hp.get("use_cuda_graphs", False):
            context.experimental.enable_cuda_graphs()


class CIFARTrial(_ClassifierTrial):
    shape, num_classes = (3, 32, 32), 10
    train_size, val_size = 8192, 2048

    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        self.model = context.wrap_model(CifarCNN(float(hp.get("layer1_dropout", 0.25)), float(hp.get("layer2_dropout", 0.25)),
                                                 float(hp.get("layer3_dropout", 0.5))))
        self.optimizer = context.wrap_optimizer(torch.optim.RMSprop(
            self.model.parameters(), lr=float(hp.get("learning_rate", 1e-4)), weight_decay=float(hp.get("learning_rate_decay", 1e-6)),
            alpha=0.9))
        if hp.get("use_cuda_graphs", False):      # small networks are launch-bound in eager mode
            context.experimental.enable_cuda_graphs()


class ResNet50Trial(_ClassifierTrial):
    shape, num_classes = (3, 224, 224), 1000
    train_size, val_size = 1024, 256
    uint8_input, prefetch = True, 2

    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        res = int(hp.get("image_size", 224))
        self.shape = (3, res, res)
        self.num_classes = int(hp.get("num_classes", 1000))
        arch = str(hp.get("arch", "resnet50"))
        net = {"resnet50": resnet50, "resnet18": resnet18}[arch](self.num_classes)
        self.model = context.wrap_model(net)
        self.optimizer = context.wrap_optimizer(torch.optim.SGD(
            self.model.parameters(), lr=float(hp.get("learning_rate", 0.1)), momentum=float(hp.get("momentum", 0.9)),
            weight_decay=float(hp.get("weight_decay", 1e-4))))
        self.lr_scheduler = context.wrap_lr_scheduler(
            torch.optim.lr_scheduler.StepLR(self.optimizer, step_size=int(hp.get("lr_step_epochs", 30)), gamma=0.1),
            dp.LRScheduler.StepMode.STEP_EVERY_EPOCH)
        if hp.get("overlap_grad_allreduce", False):
            context.experimental.enable_grad_overlap()
        if hp.get("use_cuda_graphs", False):
            context.experimental.enable_cuda_graphs()


def _bce_logits(logits: torch.Tensor, target: float) -> torch.Tensor:
    return nn.functional.binary_cross_entropy_with_logits(logits.float(), torch.full_like(logits, target, dtype=torch.float32))


class _GANBase(dp.PyTorchTrial):
    """Alternating generator / discriminator updates with two wrapped models and two wrapped optimizers."""

    latent_dim = 128
    shape: Tuple[int, int, int] = (1, 28, 28)

    def _noise(self, n: int) -> torch.Tensor:
        # drawn on the training device: no host round trip, and capturable in a CUDA graph
        return torch.randn(n, self.latent_dim, device=self.context.device)

    def build_training_data_loader(self) -> dp.DataLoader:
        # a batch is 6 MB of fp32 pixels gathered on the host: assemble the next ones in a background thread, or a 3 ms step
        # waits for it (and eight ranks on one host wait longer)
        return _loader(self.context, _dataset(self.context, "train", 4096, self.shape, 10), True, prefetch=2)

    def build_validation_data_loader(self) -> dp.DataLoader:
        return _loader(self.context, _dataset(self.context, "validation", 512, self.shape, 10), False)

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        imgs, _ = batch
        n = imgs.shape[0]
        # generator update: only the generator's parameters receive gradients
        self.generator.requires_grad_(True)
        self.discriminator.requires_grad_(False)
        fake = self.generator(self._noise(n))
        g_loss = _bce_logits(self.discriminator(fake), 1.0)
        self.context.backward(g_loss)
        self.context.step_optimizer(self.opt_g)
        # discriminator update on a real batch and the (detached) fake batch
        self.generator.requires_grad_(False)
        self.discriminator.requires_grad_(True)
        real_loss = _bce_logits(self.discriminator(imgs), 1.0)
        fake_loss = _bce_logits(self.discriminator(fake.detach()), 0.0)
        d_loss = (real_loss + fake_loss) / 2
        self.context.backward(d_loss)
        self.context.step_optimizer(self.opt_d)
        return {"loss": d_loss.detach(), "g_loss": g_loss.detach(), "d_loss": d_loss.detach()}

    def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
        imgs, _ = batch
        return {"loss": _bce_logits(self.discriminator(imgs), 1.0)}

    def _optimizers(self, hp) -> None:
        lr, b1, b2 = float(hp.get("lr", 2e-4)), float(hp.get("b1", 0.5)), float(hp.get("b2", 0.999))
        self.opt_g = self.context.wrap_optimizer(torch.optim.Adam(self.generator.parameters(), lr=lr, betas=(b1, b2)))
        self.opt_d = self.context.wrap_optimizer(torch.optim.Adam(self.discriminator.parameters(), lr=lr, betas=(b1, b2)))
        self.lr_g = self.context.wrap_lr_scheduler(
            torch.optim.lr_scheduler.LambdaLR(self.opt_g, lr_lambda=lambda epoch: 0.95 ** epoch),
            dp.LRScheduler.StepMode.STEP_EVERY_EPOCH)
        if hp.get("use_cuda_graphs", False):
            # both updates of a batch (generator, then discriminator) are captured into one graph: the requires_grad_
            # toggling is a fixed pattern and the noise is drawn on the device
            self.context.experimental.enable_cuda_graphs()


class GANTrial(_GANBase):
    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        self.latent_dim = int(hp.get("latent_dim", 128))
        self.shape = (1, 28, 28)
        self.generator = context.wrap_model(MLPGenerator(self.latent_dim, self.shape))
        self.discriminator = context.wrap_model(MLPDiscriminator(self.shape))
        self._optimizers(hp)


class DCGANTrial(_GANBase):
    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        self.latent_dim = int(hp.get("latent_dim", 128))
        nc = int(hp.get("channels", 3))
        self.shape = (nc, 64, 64)
        g = DCGANGenerator(self.latent_dim, int(hp.get("ngf", 64)), nc)
        d = DCGANDiscriminator(int(hp.get("ndf", 64)), nc)
        g.apply(dcgan_init)
        d.apply(dcgan_init)
        self.generator = context.wrap_model(g)
        self.discriminator = context.wrap_model(d)
        self._optimizers(hp)
