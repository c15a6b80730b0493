"""Model-zoo trials on CPU (fp32 torch fallback of every module): each example trains, validates and — for the
classifiers — learns on its synthetic data; conv modules interchange state_dicts with torch.nn. The CUDA kernels
behind the same modules are checked in tests/test_gpu_conv.py."""
import pytest
import torch
import torch.nn as nn

from determined_b200.models import vision_trials as VT
from determined_b200.models.resnet import resnet50
from determined_b200.ops import conv as C
from tests.harness_utils import TrainAndValidate, make_controller


def _run(trial_cls, hp, steps=3, unit=4, data=None):
    tv = TrainAndValidate()
    make_controller(trial_cls, hp, tv.send(steps=steps, validation_freq=steps, scheduling_unit=unit),
                    exp_config={"data": data or {"train_size": 64, "validation_size": 32}}).run()
    return tv


def test_mnist_trial_learns():
    tv = _run(VT.MNistTrial, {"global_batch_size": 16, "learning_rate": 1.0, "dropout1": 0.0, "dropout2": 0.0}, steps=4)
    losses = [m["loss"] for m in tv.training_metrics]
    assert len(losses) == 16 and losses[-1] < losses[0]
    assert set(tv.validation_metrics[-1]) == {"validation_loss", "accuracy", "validation_error"}


def test_cifar_trial_runs():
    tv = _run(VT.CIFARTrial, {"global_batch_size": 8, "learning_rate": 1e-3}, steps=2, unit=2)
    assert len(tv.training_metrics) == 4 and 0.0 <= tv.validation_metrics[-1]["validation_error"] <= 1.0


def test_resnet_trial_runs_small():
    hp = {"global_batch_size": 4, "arch": "resnet18", "image_size": 32, "num_classes": 16, "learning_rate": 0.01}
    tv = _run(VT.ResNet50Trial, hp, steps=1, unit=2, data={"train_size": 16, "validation_size": 8})
    assert len(tv.training_metrics) == 2 and all(torch.isfinite(torch.tensor(m["loss"])) for m in tv.training_metrics)


@pytest.mark.parametrize("trial_cls,hp", [
    (VT.GANTrial, {"global_batch_size": 8, "latent_dim": 16}),
    (VT.DCGANTrial, {"global_batch_size": 4, "latent_dim": 16, "ngf": 8, "ndf": 8, "channels": 3}),
])
def test_gan_trials_two_models_two_optimizers(trial_cls, hp):
    tv = _run(trial_cls, hp, steps=1, unit=2, data={"train_size": 16, "validation_size": 8})
    assert set(tv.training_metrics[0]) == {"loss", "g_loss", "d_loss"}
    assert "loss" in tv.validation_metrics[-1]


def test_conv_state_dict_interchange_with_torch():
    mine = nn.Sequential(C.Conv2d(3, 8, 3, padding=1), C.BatchNorm2d(8, activation="relu"), C.ConvTranspose2d(8, 4, 4, 2, 1))
    ref = nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.BatchNorm2d(8), nn.ConvTranspose2d(8, 4, 4, 2, 1))
    ref.load_state_dict(mine.state_dict())
    x = torch.randn(2, 3, 8, 8)
    y_ref = ref[2](torch.relu(ref[1](ref[0](x))))
    assert torch.allclose(mine(x), y_ref, atol=1e-5)
    mine.load_state_dict(ref.state_dict())
    assert mine[0].weight.shape == (8, 3, 3, 3) and mine[2].weight.shape == (8, 4, 4, 4)


def test_resnet50_matches_torchvision_keys_and_shapes():
    net = resnet50(10)
    sd = net.state_dict()
    assert sd["conv1.weight"].shape == (64, 3, 7, 7)
    assert sd["layer1.0.downsample.0.weight"].shape == (256, 64, 1, 1)
    assert sd["layer4.2.conv2.weight"].shape == (512, 512, 3, 3)
    assert sd["fc.weight"].shape == (10, 2048)
    assert sum(p.numel() for p in net.parameters()) == 23528522   # torchvision resnet50 with a 10-way head
