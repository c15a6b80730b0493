"""Small vision / GAN networks of the reference's example zoo, rebuilt on the sm_100a modules.

* :class:`MnistCNN` — the tutorial network (2 conv + 2 fc, dropout), reference
  ``examples/tutorials/mnist_pytorch/model_def.py:34-53``.
* :class:`CifarCNN` — the 4-conv CIFAR-10 network, reference ``examples/computer_vision/cifar10_pytorch/model_def.py:49-74``.
* :class:`MLPGenerator` / :class:`MLPDiscriminator` — the MNIST GAN, reference
  ``examples/gan/gan_mnist_pytorch/model_def.py:30-72``.
* :class:`DCGANGenerator` / :class:`DCGANDiscriminator` — the standard 64x64 DCGAN pair; the reference has only a
  TF-Keras DCGAN (``examples/gan/dcgan_tf_keras``), BASELINE config #4 asks for a PyTorchTrial one.

All of them return LOGITS where the reference returns (log-)probabilities; the matching losses live in the trials.
"""
from __future__ import annotations

from typing import Tuple

import torch
import torch.nn as nn

from ..ops.conv import Activation, BatchNorm2d, Conv2d, ConvTranspose2d, Dropout, MaxPool2d
from ..ops.modules import Linear


class _Act1d(nn.Module):
    """Activation on [N, F] features (bf16 on CUDA → kernel, otherwise torch)."""

    def __init__(self, kind: str, slope: float = 0.01) -> None:
        super().__init__()
        self.act = Activation(kind, slope)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.act(x)


class MnistCNN(nn.Module):
    def __init__(self, n_filters1: int = 32, n_filters2: int = 64, dropout1: float = 0.25, dropout2: float = 0.5) -> None:
        super().__init__()
        self.conv1 = Conv2d(1, n_filters1, 3, activation="relu")
        self.conv2 = Conv2d(n_filters1, n_filters2, 3, activation="relu")
        self.pool = MaxPool2d(2)
        self.drop1 = Dropout(dropout1)
        self.fc1 = Linear(144 * n_filters2, 128)
        self.relu = _Act1d("relu")
        self.drop2 = Dropout(dropout2)
        self.fc2 = Linear(128, 16)   # 10 classes padded to 16 outputs so the GEMM's TMA rows stay 16-byte aligned
        self.num_classes = 10

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.drop1(self.pool(self.conv2(self.conv1(x))))
        x = self.drop2(self.relu(self.fc1(x.flatten(1))))
        return self.fc2(x)[:, :self.num_classes]


class CifarCNN(nn.Module):
    def __init__(self, d1: float = 0.25, d2: float = 0.25, d3: float = 0.5, num_classes: int = 10) -> None:
        super().__init__()
        self.features = nn.Sequential(
            Conv2d(3, 32, 3, activation="relu"), Conv2d(32, 32, 3, activation="relu"), MaxPool2d(2), Dropout(d1),
            Conv2d(32, 64, 3, padding=1, activation="relu"), Conv2d(64, 64, 3, activation="relu"), MaxPool2d(2), Dropout(d2))
        self.fc1 = Linear(2304, 512)
        self.relu = _Act1d("relu")
        self.drop = Dropout(d3)
        self.fc2 = Linear(512, 16)
        self.num_classes = num_classes

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.features(x).flatten(1)
        return self.fc2(self.drop(self.relu(self.fc1(x))))[:, :self.num_classes]


class MLPGenerator(nn.Module):
    def __init__(self, latent_dim: int = 128, img_shape: Tuple[int, int, int] = (1, 28, 28)) -> None:
        super().__init__()
        self.img_shape = img_shape
        out = img_shape[0] * img_shape[1] * img_shape[2]
        self.net = nn.Sequential(Linear(latent_dim, 128), _Act1d("leaky_relu", 0.2), Linear(128, 256), _Act1d("leaky_relu", 0.2),
                                 Linear(256, 512), _Act1d("leaky_relu", 0.2), Linear(512, 1024), _Act1d("leaky_relu", 0.2),
                                 Linear(1024, out), _Act1d("tanh"))

    def forward(self, z: torch.Tensor) -> torch.Tensor:
        return self.net(z).view(z.shape[0], *self.img_shape)


class MLPDiscriminator(nn.Module):
    def __init__(self, img_shape: Tuple[int, int, int] = (1, 28, 28)) -> None:
        super().__init__()
        n_in = img_shape[0] * img_shape[1] * img_shape[2]
        self.net = nn.Sequential(Linear(n_in, 512), _Act1d("leaky_relu", 0.2), Linear(512, 256), _Act1d("leaky_relu", 0.2),
                                 Linear(256, 8))   # one logit, padded to 8 outputs for TMA alignment

    def forward(self, img: torch.Tensor) -> torch.Tensor:
        return self.net(img.reshape(img.shape[0], -1))[:, :1]


class DCGANGenerator(nn.Module):
    """z [N, nz, 1, 1] → image [N, nc, 64, 64] in (-1, 1)."""

    def __init__(self, nz: int = 128, ngf: int = 64, nc: int = 3) -> None:
        super().__init__()
        self.nz = nz
        self.net = nn.Sequential(
            ConvTranspose2d(nz, ngf * 8, 4, 1, 0, bias=False), BatchNorm2d(ngf * 8, activation="relu"),
            ConvTranspose2d(ngf * 8, ngf * 4, 4, 2, 1, bias=False), BatchNorm2d(ngf * 4, activation="relu"),
            ConvTranspose2d(ngf * 4, ngf * 2, 4, 2, 1, bias=False), BatchNorm2d(ngf * 2, activation="relu"),
            ConvTranspose2d(ngf * 2, ngf, 4, 2, 1, bias=False), BatchNorm2d(ngf, activation="relu"),
            ConvTranspose2d(ngf, nc, 4, 2, 1, bias=False), Activation("tanh"))

    def forward(self, z: torch.Tensor) -> torch.Tensor:
        return self.net(z.view(z.shape[0], self.nz, 1, 1))


class DCGANDiscriminator(nn.Module):
    """image [N, nc, 64, 64] → one logit per image."""

    def __init__(self, ndf: int = 64, nc: int = 3) -> None:
        super().__init__()
        self.net = nn.Sequential(
            Conv2d(nc, ndf, 4, 2, 1, bias=False), Activation("leaky_relu", 0.2),
            Conv2d(ndf, ndf * 2, 4, 2, 1, bias=False), BatchNorm2d(ndf * 2, activation="leaky_relu", negative_slope=0.2),
            Conv2d(ndf * 2, ndf * 4, 4, 2, 1, bias=False), BatchNorm2d(ndf * 4, activation="leaky_relu", negative_slope=0.2),
            Conv2d(ndf * 4, ndf * 8, 4, 2, 1, bias=False), BatchNorm2d(ndf * 8, activation="leaky_relu", negative_slope=0.2),
            Conv2d(ndf * 8, 1, 4, 1, 0, bias=False))

    def forward(self, img: torch.Tensor) -> torch.Tensor:
        return self.net(img).reshape(img.shape[0], 1)


def dcgan_init(m: nn.Module) -> None:
    """N(0, 0.02) weights / N(1, 0.02) BatchNorm scales, as in the DCGAN paper."""
    if isinstance(m, (Conv2d, ConvTranspose2d)):
        nn.init.normal_(m.weight, 0.0, 0.02)
    elif isinstance(m, BatchNorm2d):
        nn.init.normal_(m.weight, 1.0, 0.02)
        nn.init.zeros_(m.bias)
