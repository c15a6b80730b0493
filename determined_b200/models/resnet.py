"""ResNet (bottleneck) image classifiers on the sm_100a convolution stack.

Same graph and the same ``state_dict`` keys as torchvision's ``resnet50`` (v1.5: the stride sits on the 3x3
convolution), so ImageNet checkpoints load unchanged. The reference ships no ResNet-50 example — BASELINE config #2
("ResNet-50 PyTorchTrial, synthetic 224x224 ImageNet-shape batches") had to be authored new; the
closest reference model definitions are ``examples/computer_vision/cifar10_pytorch/model_def.py:49-74`` and the
torchvision backbones pulled in by ``examples/computer_vision/fasterrcnn_coco_pytorch``.

Per bottleneck block: 1x1 convolutions are plain tcgen05 GEMMs on the NHWC array, the 3x3 is the implicit-GEMM kernel
(or the gather path when strided), every BatchNorm fuses its ReLU and the last one also the residual add.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.nn as nn

from ..ops.conv import BatchNorm2d, Conv2d, GlobalAvgPool2d, MaxPool2d, conv_bn
from ..ops.modules import Linear


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None) -> None:
        super().__init__()
        self.conv1 = Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = BatchNorm2d(planes, activation="relu")
        self.conv2 = Conv2d(planes, planes, 3, stride=stride, padding=1, bias=False)
        self.bn2 = BatchNorm2d(planes, activation="relu")
        self.conv3 = Conv2d(planes, planes * self.expansion, 1, bias=False)
        self.bn3 = BatchNorm2d(planes * self.expansion, activation="relu")  # relu(bn(x) + identity) in one kernel
        self.downsample = downsample

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.downsample is None:
            # the skip connection leaves through conv1's autograd node: its gradient is added in conv1's dgrad epilogue
            out, identity = conv_bn(self.conv1, self.bn1, x, passthrough=True)
        else:
            identity = conv_bn(self.downsample[0], self.downsample[1], x)
            out = conv_bn(self.conv1, self.bn1, x)
        out = conv_bn(self.conv2, self.bn2, out)
        return conv_bn(self.conv3, self.bn3, out, identity)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None) -> None:
        super().__init__()
        self.conv1 = Conv2d(inplanes, planes, 3, stride=stride, padding=1, bias=False)
        self.bn1 = BatchNorm2d(planes, activation="relu")
        self.conv2 = Conv2d(planes, planes, 3, padding=1, bias=False)
        self.bn2 = BatchNorm2d(planes, activation="relu")
        self.downsample = downsample

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        identity = x if self.downsample is None else conv_bn(self.downsample[0], self.downsample[1], x)
        return conv_bn(self.conv2, self.bn2, conv_bn(self.conv1, self.bn1, x), identity)


class ResNet(nn.Module):
    def __init__(self, block, layers: List[int], num_classes: int = 1000, in_channels: int = 3) -> None:
        super().__init__()
        self.inplanes = 64
        self.conv1 = Conv2d(in_channels, 64, 7, stride=2, padding=3, bias=False)
        self.bn1 = BatchNorm2d(64, activation="relu")
        self.maxpool = MaxPool2d(3, 2, 1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.layer3 = self._make_layer(block, 256, layers[2], stride=2)
        self.layer4 = self._make_layer(block, 512, layers[3], stride=2)
        self.avgpool = GlobalAvgPool2d()
        self.fc = Linear(512 * block.expansion, num_classes)
        for m in self.modules():
            if isinstance(m, Conv2d):
                nn.init.kaiming_normal_(m.weight.permute(0, 3, 1, 2), mode="fan_out", nonlinearity="relu")

    def _make_layer(self, block, planes: int, blocks: int, stride: int = 1) -> nn.Sequential:
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(Conv2d(self.inplanes, planes * block.expansion, 1, stride=stride, bias=False),
                                       BatchNorm2d(planes * block.expansion))
        layers = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        layers += [block(self.inplanes, planes) for _ in range(1, blocks)]
        return nn.Sequential(*layers)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.maxpool(conv_bn(self.conv1, self.bn1, x))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        return self.fc(self.avgpool(x))


def resnet18(num_classes: int = 1000) -> ResNet:
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes)


def resnet50(num_classes: int = 1000) -> ResNet:
    return ResNet(Bottleneck, [3, 4, 6, 3], num_classes)
