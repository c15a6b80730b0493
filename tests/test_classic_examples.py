"""The PyTorch counterparts of the reference's TensorFlow-only examples (Fashion-MNIST tutorial, Iris, U-Net segmentation, MNIST
through the data layer, boosted trees on Titanic): each one trains through a real controller, learns its synthetic task, and
checkpoints / restores; the booster is also checked against a plain second-order reference of one split."""
import pathlib
import subprocess
import sys

import pytest
import torch

from determined_b200.models import boosted_trees as bt
from determined_b200.models import classic_trials as ct
from tests import harness_utils as utils

ROOT = pathlib.Path(__file__).resolve().parent.parent


def run(trial_class, hp, steps, scheduling_unit, tmp_path=None, exp_config=None, load_path=None, batches_done=0):
    tv = utils.TrainAndValidate()
    ckpt = tmp_path / "ckpt" if tmp_path is not None else None
    controller = utils.make_controller(trial_class, hp, tv.send(steps, validation_freq=steps, scheduling_unit=scheduling_unit,
                                                                checkpoint_dir=ckpt, batches_done=batches_done),
                                       exp_config=exp_config, load_path=load_path, initial_batches=batches_done)
    controller.run()
    return tv, ckpt


def test_fashion_mnist_learns():
    cfg = {"data": {"train_size": 1024, "validation_size": 256}, "searcher": {"name": "single", "metric": "val_accuracy", "max_length": {"batches": 100}}}
    tv, _ = run(ct.FashionMNistTrial, {"global_batch_size": 32, "learning_rate": 1e-3, "dense1": 64, "dropout": 0.1}, 4, 16, exp_config=cfg)
    assert tv.avg_metrics[-1]["loss"] < tv.avg_metrics[0]["loss"]
    assert tv.validation_metrics[-1]["val_accuracy"] > 0.8


def test_iris_learns_and_decays_its_learning_rate():
    cfg = {"data": {"train_size": 120, "validation_size": 30}, "searcher": {"name": "single", "metric": "val_categorical_accuracy", "max_length": {"batches": 400}}}
    hp = {"global_batch_size": 30, "learning_rate": 1e-2, "learning_rate_decay": 1e-2, "layer1_dense_size": 16, "layer2_dense_size": 16}
    tv = utils.TrainAndValidate()
    controller = utils.make_controller(ct.IrisTrial, hp, tv.send(8, validation_freq=8, scheduling_unit=25), exp_config=cfg)
    controller.run()
    assert tv.validation_metrics[-1]["val_categorical_accuracy"] >= 0.9
    lr = controller.trial.optimizer.param_groups[0]["lr"]
    assert lr == pytest.approx(1e-2 / (1 + 1e-2 * 200), rel=1e-6)          # Keras-style decay: lr / (1 + decay * iterations)


def test_unet_segments_and_restores(tmp_path):
    cfg = {"data": {"train_size": 96, "validation_size": 32}, "searcher": {"name": "single", "metric": "val_accuracy", "max_length": {"batches": 100}}}
    hp = {"global_batch_size": 16, "learning_rate": 3e-3, "image_size": 32, "width": 8}
    net = ct.UNet(3, 8)
    assert net(torch.zeros(2, 3, 32, 32)).shape == (2, 3, 32, 32) and net(torch.zeros(1, 3, 128, 128)).shape == (1, 3, 128, 128)
    tv, ckpt = run(ct.UNetTrial, hp, 3, 6, tmp_path, cfg)
    assert tv.avg_metrics[-1]["loss"] < tv.avg_metrics[0]["loss"]
    assert tv.validation_metrics[-1]["val_accuracy"] > 0.7 and 0.0 <= tv.validation_metrics[-1]["val_pet_iou"] <= 1.0
    # a restored trial validates to the same numbers without training (eval mode: batch-norm statistics come from the checkpoint)
    tv2 = utils.TrainAndValidate()

    def only_validate():
        from determined_b200 import workload as wl

        interceptor = wl.WorkloadResponseInterceptor()
        yield from interceptor.send(wl.validation_workload(3, total_batches_processed=18), [])
        tv2.validation_metrics.append(interceptor.metrics_result()["metrics"]["validation_metrics"])
        yield from interceptor.send(wl.terminate_workload(), [])

    utils.make_controller(ct.UNetTrial, hp, only_validate(), exp_config=cfg, load_path=ckpt, initial_batches=18).run()
    assert tv2.validation_metrics[-1]["val_loss"] == pytest.approx(tv.validation_metrics[-1]["val_loss"], rel=1e-5)


def test_data_layer_mnist_builds_once_and_reuses_the_cache(tmp_path, monkeypatch):
    calls = []
    real = ct.synthetic_images
    monkeypatch.setattr(ct, "synthetic_images", lambda *a, **k: (calls.append(a), real(*a, **k))[1])
    cfg = {"data": {"train_size": 256, "validation_size": 64, "dataset_version": "7"},
           "data_layer": {"type": "shared_fs", "container_storage_path": str(tmp_path / "cache")},
           "searcher": {"name": "single", "metric": "val_accuracy", "max_length": {"batches": 100}}}
    hp = {"global_batch_size": 32, "learning_rate": 1e-3, "dense1": 32, "dropout": 0.0}
    tv, _ = run(ct.DataLayerMNistTrial, hp, 2, 8, exp_config=cfg)
    assert len(calls) == 2 and tv.validation_metrics[-1]["val_accuracy"] > 0.5           # train + validation built
    tv, _ = run(ct.DataLayerMNistTrial, hp, 1, 8, exp_config=cfg)
    assert len(calls) == 2                                                              # second trial: served from the cache
    cfg["data"]["dataset_version"] = "8"
    run(ct.DataLayerMNistTrial, hp, 1, 8, exp_config=cfg)
    assert len(calls) == 4                                                              # a new version is built again
    assert any((tmp_path / "cache").rglob("*"))


# ---- boosted trees ---------------------------------------------------------------------------------------------------------
def test_first_split_matches_a_brute_force_second_order_search():
    x, y = bt.synthetic_titanic(400, 3)
    m = bt.GradientBoostedClassifier(max_depth=1, learning_rate=1.0, l2=1.0, n_bins=16)
    m.fit_bins(x)
    xb = m.quantise(x)
    margin = torch.zeros(400)
    tree, new_margin = m.boost(xb, y, margin)
    g, h = torch.sigmoid(margin) - y, torch.full((400,), 0.25)
    score = lambda G, H: G * G / (H + 1.0)  # noqa: E731
    best, best_gain = None, -1.0
    for f in range(xb.shape[1]):
        for b in range(15):
            left = xb[:, f] <= b
            gain = float(score(g[left].sum(), h[left].sum()) + score(g[~left].sum(), h[~left].sum()) - score(g.sum(), h.sum()))
            if gain > best_gain + 1e-9:
                best, best_gain = (f, b), gain
    assert (int(tree.features[0]), int(tree.bins[0])) == best
    left = xb[:, best[0]] <= best[1]
    want = torch.where(left, -g[left].sum() / (h[left].sum() + 1.0), -g[~left].sum() / (h[~left].sum() + 1.0))
    assert torch.allclose(new_margin, want, atol=1e-5)
    assert int(tree.features[0]) == 0                      # and on this data the first question is the passenger's sex


def test_regularisation_and_minimum_node_weight():
    x, y = bt.synthetic_titanic(300, 4)
    big = bt.GradientBoostedClassifier(max_depth=3, learning_rate=1.0, l2=0.0)
    small = bt.GradientBoostedClassifier(max_depth=3, learning_rate=1.0, l2=50.0)
    sparse = bt.GradientBoostedClassifier(max_depth=3, learning_rate=1.0, l1=1e6)
    for m in (big, small, sparse):
        m.fit_bins(x)
    xb = big.quantise(x)
    t_big, _ = big.boost(xb, y, torch.zeros(300))
    t_small, _ = small.boost(xb, y, torch.zeros(300))
    t_sparse, _ = sparse.boost(xb, y, torch.zeros(300))
    assert t_small.leaves.abs().max() < t_big.leaves.abs().max()          # L2 shrinks the leaf values
    assert float(t_sparse.leaves.abs().max()) == 0.0                        # an L1 penalty above every gradient sum zeroes them
    heavy = bt.GradientBoostedClassifier(max_depth=2, learning_rate=1.0, min_node_weight=1e9)
    heavy.fit_bins(x)
    _, margin = heavy.boost(xb, y, torch.zeros(300))
    assert torch.isfinite(margin).all()                                     # no admissible split: still a valid (constant) tree


def test_gbt_trial_learns_stops_growing_at_n_trees_and_resumes(tmp_path):
    cfg = {"data": {"train_size": 600, "validation_size": 200}, "searcher": {"name": "single", "metric": "accuracy", "smaller_is_better": False,
                                                                                 "max_length": {"batches": 40}}}
    hp = {"global_batch_size": 600, "n_trees": 12, "max_depth": 4, "learning_rate": 0.3, "l2_regularization": 1.0}
    tv, ckpt = run(bt.GBTTitanicTrial, hp, 2, 5, tmp_path, cfg)
    assert tv.training_metrics[-1]["loss"] < tv.training_metrics[0]["loss"]
    v = tv.validation_metrics[-1]
    assert v["n_trees"] == 10 and v["accuracy"] > 0.7
    # resume from the checkpoint: two more rounds complete the ensemble, later rounds add nothing
    tv2 = utils.TrainAndValidate()
    controller = utils.make_controller(bt.GBTTitanicTrial, hp, tv2.send(1, validation_freq=1, scheduling_unit=5, batches_done=10), exp_config=cfg,
                                       load_path=ckpt, initial_batches=10)
    controller.run()
    assert tv2.validation_metrics[-1]["n_trees"] == 12
    assert tv2.training_metrics[1]["loss"] == pytest.approx(tv2.training_metrics[-1]["loss"])     # rounds 3..5 changed nothing
    assert tv2.training_metrics[0]["loss"] < tv.training_metrics[-1]["loss"]                       # and it continued, not restarted


@pytest.mark.parametrize("example", ["tutorials/fashion_mnist_pytorch", "computer_vision/iris_pytorch", "decision_trees/gbt_titanic_pytorch",
                                     "data_layer/data_layer_mnist_pytorch"])
def test_example_local_test_mode(example):
    d = ROOT / "examples" / example
    out = subprocess.run([sys.executable, "-m", "determined_b200.cli", "experiment", "create", str(d / "const.yaml"), str(d), "--local", "--test"],
                         cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "Model definition test succeeded" in out.stdout + out.stderr
