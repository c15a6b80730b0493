"""GAEA weight-sharing search (reference: examples/nas/gaea_pytorch/search): simplex-constrained architecture parameters,
the exponentiated-gradient update, genotype extraction and its round trip into the evaluation trial's hyperparameters."""
import torch

from determined_b200.models import darts as D
from determined_b200.models import gaea_search as G
from tests.harness_utils import TrainAndValidate, make_controller


def test_exponentiated_gradient_stays_on_the_simplex_and_sparsifies():
    p = torch.nn.Parameter(torch.full((3, 8), 1 / 8))
    opt = G.ExponentiatedGradient([p], lr=0.5)
    grad = torch.zeros(3, 8)
    grad[:, 2] = -1.0                       # operation 2 keeps reducing the loss
    for _ in range(20):
        p.grad = grad.clone()
        opt.step()
    assert torch.allclose(p.sum(-1), torch.ones(3), atol=1e-6) and (p > 0).all()
    assert (p.argmax(-1) == 2).all() and p[0, 2] > 0.99


def test_supernet_forward_backward_and_genotype():
    torch.manual_seed(0)
    net = G.SearchNetwork(8, 10, layers=3)
    assert net.alphas_normal.shape == (14, 8) and len(net.weight_parameters()) + 2 == len(list(net.parameters()))
    logits = net(torch.randn(4, 3, 32, 32))
    assert logits.shape == (4, 10)
    logits.sum().backward()
    assert net.alphas_normal.grad.abs().sum() > 0 and net.alphas_reduce.grad.abs().sum() > 0
    assert net.alphas_normal.grad[:, 0].abs().sum() == 0          # "none" computes nothing
    with torch.no_grad():
        net.alphas_normal.fill_(0.01)
        net.alphas_normal[0, 4] = 0.9      # edge 0 (node 1 <- state 0): sep_conv_3x3
        net.alphas_normal[1, 1] = 0.8      # edge 1 (node 1 <- state 1): max_pool_3x3
        net.alphas_normal[4, 0] = 0.99     # a strong "none" must not make an edge win
    g = net.genotype()
    assert g.normal[0] == ("sep_conv_3x3", 0) and g.normal[1] == ("max_pool_3x3", 1) and len(g.normal) == 8
    hp = G.genotype_to_hparams(g)
    assert D.genotype_from_hparams(hp) == g
    D.DartsNetworkCIFAR(8, 10, layers=3, auxiliary=False, genotype=g)(torch.randn(2, 3, 32, 32))


def test_search_trial_updates_weights_and_architecture():
    hp = {"global_batch_size": 16, "learning_rate": 0.05, "arch_learning_rate": 0.5, "init_channels": 8, "layers": 3, "train_epochs": 5}
    tv = TrainAndValidate()
    c = make_controller(G.GAEASearchTrial, hp, tv.send(steps=2, validation_freq=2, scheduling_unit=6),
                        exp_config={"data": {"train_size": 64, "validation_size": 32},
                                    "searcher": {"name": "single", "metric": "top1_accuracy", "smaller_is_better": False, "max_length": {"batches": 100}}})
    c.run()
    assert len(tv.training_metrics) == 12 and {"loss", "arch_loss", "top1_accuracy"} <= set(tv.training_metrics[0])
    net = c.trial.model
    a = net.alphas_normal.detach()
    assert torch.allclose(a.sum(-1), torch.ones(14), atol=1e-5) and (a - 1 / 8).abs().max() > 1e-4   # moved, still on the simplex
    assert all(p.requires_grad for p in net.weight_parameters())
    assert {"loss", "top1_accuracy", "top5_accuracy", "arch_entropy"} == set(tv.validation_metrics[-1])
