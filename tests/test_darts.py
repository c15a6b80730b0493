"""DARTS cell network and its building blocks (reference: examples/hp_search_benchmarks/darts_cifar10_pytorch): depthwise /
windowed-average modules against torch, genotype decoding, and the trial learning through the controller."""
import pytest
import torch
import torch.nn as nn

from determined_b200.models import darts as D
from determined_b200.models.darts_trial import DARTSCNNTrial, cutout_
from determined_b200.ops.conv import AvgPool2d, DepthwiseConv2d
from tests.harness_utils import TrainAndValidate, make_controller


def test_depthwise_state_dict_interchange_and_cpu_math():
    mine = DepthwiseConv2d(16, 3, stride=2, padding=2, dilation=2, bias=True)
    ref = nn.Conv2d(16, 16, 3, 2, 2, 2, groups=16)
    assert mine.state_dict()["weight"].shape == ref.weight.shape
    mine.load_state_dict(ref.state_dict())
    x = torch.randn(3, 16, 13, 11)
    assert torch.allclose(mine(x), ref(x), atol=1e-6)
    ap, rp = AvgPool2d(3, 2, 1, count_include_pad=False), nn.AvgPool2d(3, 2, 1, count_include_pad=False)
    assert torch.allclose(ap(x), rp(x))


def test_genotype_decoding():
    g = D.genotype_from_hparams({})
    assert g == D.DARTS_V2
    g = D.genotype_from_hparams({"normal_node1_edge1": 1, "normal_node1_edge1_op": "dil_conv_5x5", "reduce_node4_edge2": 4,
                                 "reduce_node4_edge2_op": "avg_pool_3x3"})
    assert g.normal[0] == ("dil_conv_5x5", 1) and g.reduce[7] == ("avg_pool_3x3", 4) and g.normal[1:] == D.DARTS_V2.normal[1:]
    with pytest.raises(ValueError):
        D.genotype_from_hparams({"normal_node1_edge1": 2})          # node 1 can only read states 0 and 1
    with pytest.raises(ValueError):
        D.genotype_from_hparams({"normal_node2_edge1_op": "conv_7x7"})


def test_network_shapes_and_every_primitive():
    hp = {}
    ops = D.PRIMITIVES[1:]
    k = 0
    for cell in ("normal", "reduce"):
        for i in range(1, 5):
            for j in range(1, 3):
                hp[f"{cell}_node{i}_edge{j}_op"] = ops[k % len(ops)]
                hp[f"{cell}_node{i}_edge{j}"] = k % (i + 1)
                k += 1
    net = D.DartsNetworkCIFAR(8, 10, layers=6, auxiliary=True, genotype=D.genotype_from_hparams(hp))
    net.drop_path_prob = 0.3
    logits, aux = net(torch.randn(4, 3, 32, 32))
    assert logits.shape == (4, 10) and aux.shape == (4, 10)
    (logits.sum() + aux.sum()).backward()
    assert all(p.grad is not None for n, p in net.named_parameters())
    net.eval()
    assert net(torch.randn(2, 3, 32, 32))[1] is None                 # the auxiliary head only runs in training


def test_cutout_and_topk():
    x = torch.ones(8, 3, 32, 32)
    cutout_(x, 16, torch.Generator().manual_seed(0))
    holes = (x == 0).float().mean().item()
    assert 0.05 < holes < 0.3
    logits = torch.tensor([[0.1, 0.9, 0.0], [0.8, 0.1, 0.05], [0.2, 0.3, 0.5]])
    top1, top2 = D.accuracy_topk(logits, torch.tensor([1, 1, 0]), (1, 2))
    assert abs(top1.item() - 1 / 3) < 1e-6 and abs(top2.item() - 2 / 3) < 1e-6


def test_darts_trial_learns():
    hp = {"global_batch_size": 16, "learning_rate": 0.05, "init_channels": 8, "layers": 3, "auxiliary": True, "train_epochs": 5,
          "drop_path_prob": 0.1, "cutout_length": 8}
    tv = TrainAndValidate()
    make_controller(DARTSCNNTrial, hp, tv.send(steps=2, validation_freq=2, scheduling_unit=8),
                    exp_config={"data": {"train_size": 64, "validation_size": 32},
                                "searcher": {"name": "single", "metric": "top1_accuracy", "smaller_is_better": False, "max_length": {"batches": 100}}}).run()
    losses = [m["loss"] for m in tv.training_metrics]
    assert len(losses) == 16 and sum(losses[-4:]) < sum(losses[:4])
    assert {"loss", "top1_accuracy", "top5_accuracy"} == set(tv.validation_metrics[-1])
