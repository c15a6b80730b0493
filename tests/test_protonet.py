"""Prototypical-network trial (reference: examples/meta_learning/protonet_omniglot_pytorch): episode layout, the distance
form, and meta-training through the controller lifting accuracy on held-out characters well above chance."""
import torch

from determined_b200.models import protonet_trial as P
from tests.harness_utils import TrainAndValidate, make_controller


def test_episode_layout_and_distances():
    ds = P.SyntheticOmniglotTasks(5, torch.arange(40), 28, n_classes=6, n_support=2, n_query=3, seed=1)
    xs, ys, xq, yq = ds[3]
    assert xs.shape == (12, 1, 28, 28) and xq.shape == (18, 1, 28, 28)
    assert ys.tolist() == [0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5] and yq[:4].tolist() == [0, 0, 0, 1]
    a = ds[3]
    assert all(torch.equal(u, v) for u, v in zip(a, (xs, ys, xq, yq)))          # episodes are reproducible
    x, y = torch.randn(7, 16), torch.randn(5, 16)
    ref = ((x[:, None, :] - y[None, :, :]) ** 2).sum(2)
    assert torch.allclose(P.squared_distances(x, y), ref, atol=1e-4)


def test_protonet_meta_learns():
    hp = {"global_batch_size": 2, "val_batch_size": 2, "learning_rate": 2e-3, "num_classes_train": 10, "num_support_train": 1,
          "num_query_train": 3, "num_classes_val": 5, "num_support_val": 1, "num_query_val": 3, "img_resize_dim": 28,
          "hidden_dim": 16, "embedding_dim": 16, "reduce_every": 200, "lr_gamma": 0.5}
    tv = TrainAndValidate()
    make_controller(P.OmniglotProtoNetTrial, hp, tv.send(steps=3, validation_freq=3, scheduling_unit=10),
                    exp_config={"data": {"tasks_per_epoch_train": 60, "tasks_per_epoch_val": 20, "num_characters": 80},
                                "searcher": {"name": "single", "metric": "loss", "max_length": {"batches": 100}}}).run()
    losses = [m["loss"] for m in tv.training_metrics]
    assert len(losses) == 30 and sum(losses[-5:]) < sum(losses[:5])
    val = tv.validation_metrics[-1]
    assert set(val) == {"loss", "acc"} and val["acc"] > 0.5                     # 5-way chance is 0.2
