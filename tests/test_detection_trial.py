"""A stock torchvision detector inside PyTorchTrial (reference: examples/computer_vision/fasterrcnn_coco_pytorch): list
batches through a custom collate_fn, dict targets moved to the device, the IoU validation metric, and the controller
driving train / validate workloads with a model that is NOT built from this package's modules."""
import pytest
import torch

torchvision = pytest.importorskip("torchvision")

from determined_b200.models import detection_trial as DT  # noqa: E402
from tests.harness_utils import TrainAndValidate, make_controller  # noqa: E402


def test_dataset_and_collate():
    ds = DT.SyntheticPedestrians(6, 64, seed=3)
    img, tgt = ds[2]
    assert img.shape == (3, 64, 64) and tgt["boxes"].shape[1] == 4 and tgt["labels"].tolist() == [1] * len(tgt["boxes"])
    assert (tgt["boxes"][:, 2] > tgt["boxes"][:, 0]).all() and (tgt["boxes"][:, 3] <= 64).all()
    images, targets = DT.collate_fn([ds[0], ds[1], ds[2]])
    assert len(images) == 3 and isinstance(targets[0], dict)


def test_average_best_iou():
    tgt = [{"boxes": torch.tensor([[0., 0., 10., 10.], [20., 20., 30., 30.]])}]
    out = [{"boxes": torch.tensor([[0., 0., 10., 10.], [0., 0., 9., 10.], [20., 20., 30., 25.]]), "scores": torch.tensor([0.9, 0.8, 0.7])}]
    assert abs(DT.average_best_iou(out, tgt).item() - (1.0 + 0.5) / 2) < 1e-6     # the 0.8 box is suppressed; the third covers half
    assert DT.average_best_iou([{"boxes": torch.zeros(0, 4), "scores": torch.zeros(0)}], tgt).item() == 0.0


def test_detection_trial_runs_through_the_controller():
    hp = {"global_batch_size": 2, "learning_rate": 0.01, "momentum": 0.9, "weight_decay": 0.0005, "backbone": "mobilenet_v3_large_320_fpn",
          "image_size": 64, "dataset_size": 20}
    tv = TrainAndValidate()
    make_controller(DT.ObjectDetectionTrial, hp, tv.send(steps=1, validation_freq=1, scheduling_unit=3),
                    exp_config={"searcher": {"name": "single", "metric": "val_avg_iou", "smaller_is_better": False, "max_length": {"batches": 10}}}).run()
    assert len(tv.training_metrics) == 3 and all(torch.isfinite(torch.tensor(m["loss"])) for m in tv.training_metrics)
    assert {"loss", "loss_classifier", "loss_objectness"} <= set(tv.training_metrics[0])
    assert 0.0 <= tv.validation_metrics[-1]["val_avg_iou"] <= 1.0
