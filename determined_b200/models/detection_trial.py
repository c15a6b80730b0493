"""Object-detection fine-tuning with a STOCK torchvision model inside ``PyTorchTrial``.

Counterpart of the reference's ``examples/computer_vision/fasterrcnn_coco_pytorch/model_def.py`` (torchvision
``fasterrcnn_resnet50_fpn`` with a two-class box predictor on PennFudan pedestrians, SGD + StepLR per epoch, the box
regression loss as the training loss, "average best IoU per target box after NMS" as the validation metric, list-of-images
batches through a custom ``collate_fn``). It shows the other half of the platform's contract: a model does not have to be
built from the sm_100a modules — any ``torch.nn.Module`` goes through ``wrap_model`` / ``wrap_optimizer`` (flat-bucket fused
optimizer, in-kernel gradient all-reduce, checkpoints, searchers) unchanged; here the network itself runs on torch's own
kernels. Neither the COCO weights nor the PennFudan archive can be downloaded in this environment, so the detector starts
from random weights and the data are synthetic scenes: bright rectangles ("pedestrians") on textured backgrounds.

Hyperparameters: ``learning_rate``, ``momentum``, ``weight_decay``, ``global_batch_size``; optional ``backbone``
(``resnet50_fpn`` | ``mobilenet_v3_large_320_fpn``), ``image_size``, ``min_size``, ``max_size``, ``dataset_size``.
"""
from __future__ import annotations

import copy
from typing import Any, Dict, List, Tuple

import torch

from .. import pytorch as dp


class SyntheticPedestrians(torch.utils.data.Dataset):
    """``(image [3, H, W] float in [0, 1], {"boxes": [n, 4] xyxy, "labels": [n]})`` with 1-3 boxes per image."""

    def __init__(self, size: int, image_size: int = 128, seed: int = 0) -> None:
        self.size, self.hw, self.seed = size, image_size, seed

    def __len__(self) -> int:
        return self.size

    def __getitem__(self, i: int) -> Tuple[torch.Tensor, Dict[str, torch.Tensor]]:
        g = torch.Generator().manual_seed(self.seed * 100_003 + i)
        hw = self.hw
        img = torch.rand(3, hw, hw, generator=g) * 0.25
        n = int(torch.randint(1, 4, (1,), generator=g))
        boxes = []
        for _ in range(n):
            w = int(torch.randint(hw // 8, hw // 3, (1,), generator=g))
            h = int(torch.randint(hw // 4, hw // 2, (1,), generator=g))
            x0 = int(torch.randint(0, hw - w, (1,), generator=g))
            y0 = int(torch.randint(0, hw - h, (1,), generator=g))
            img[:, y0:y0 + h, x0:x0 + w] = 0.7 + 0.3 * torch.rand(3, 1, 1, generator=g)
            boxes.append([x0, y0, x0 + w, y0 + h])
        return img, {"boxes": torch.tensor(boxes, dtype=torch.float32), "labels": torch.ones(n, dtype=torch.int64)}


def collate_fn(batch: List[Tuple[torch.Tensor, Dict[str, torch.Tensor]]]):
    """Images differ in size in general, so a batch is a tuple of lists rather than a stacked tensor."""
    return tuple(zip(*batch))


def build_detector(backbone: str, num_classes: int, min_size: int, max_size: int) -> torch.nn.Module:
    from torchvision.models import detection
    from torchvision.models.detection.faster_rcnn import FastRCNNPredictor

    ctor = {"resnet50_fpn": detection.fasterrcnn_resnet50_fpn,
            "mobilenet_v3_large_320_fpn": detection.fasterrcnn_mobilenet_v3_large_320_fpn}[backbone]
    model = ctor(weights=None, weights_backbone=None, min_size=min_size, max_size=max_size)
    in_features = model.roi_heads.box_predictor.cls_score.in_features
    model.roi_heads.box_predictor = FastRCNNPredictor(in_features, num_classes)   # background + pedestrian
    return model


def average_best_iou(outputs: List[Dict[str, torch.Tensor]], targets: List[Dict[str, torch.Tensor]], nms_threshold: float = 0.1) -> torch.Tensor:
    """Per target box, the best IoU among the predictions that survive non-maximum suppression; averaged over targets."""
    from torchvision import ops

    total = torch.zeros((), device=targets[0]["boxes"].device)
    count = 0
    for out, tgt in zip(outputs, targets):
        keep = ops.nms(out["boxes"], out["scores"], nms_threshold)
        pred = out["boxes"][keep]
        count += tgt["boxes"].shape[0]
        if pred.shape[0]:
            total = total + ops.box_iou(tgt["boxes"], pred).max(dim=1).values.sum()
    return total / max(count, 1)


class ObjectDetectionTrial(dp.PyTorchTrial):
    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        self.image_size = int(hp.get("image_size", 128))
        self.dataset_size = int(hp.get("dataset_size", 170))       # PennFudan has 170 images
        model = build_detector(str(hp.get("backbone", "resnet50_fpn")), 2, int(hp.get("min_size", self.image_size)),
                               int(hp.get("max_size", self.image_size)))
        self.model = context.wrap_model(model)
        self.optimizer = context.wrap_optimizer(torch.optim.SGD(
            self.model.parameters(), lr=float(hp.get("learning_rate", 0.005)), momentum=float(hp.get("momentum", 0.9)),
            weight_decay=float(hp.get("weight_decay", 0.0005))))
        self.lr_scheduler = context.wrap_lr_scheduler(torch.optim.lr_scheduler.StepLR(self.optimizer, step_size=3, gamma=0.1),
                                                      dp.LRScheduler.StepMode.STEP_EVERY_EPOCH)

    def _loader(self, lo: int, hi: int, seed: int) -> dp.DataLoader:
        ds = torch.utils.data.Subset(SyntheticPedestrians(self.dataset_size, self.image_size, seed), range(lo, hi))
        return dp.DataLoader(ds, batch_size=self.context.get_per_slot_batch_size(), collate_fn=collate_fn, drop_last=True)

    def build_training_data_loader(self) -> dp.DataLoader:
        return self._loader(0, int(0.8 * self.dataset_size), 1)      # 80 / 20 split like the reference

    def build_validation_data_loader(self) -> dp.DataLoader:
        return self._loader(int(0.8 * self.dataset_size), self.dataset_size, 1)

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        images, targets = batch
        losses = self.model(list(images), list(targets))
        self.context.backward(losses["loss_box_reg"])                # the reference optimises the box regression term
        self.context.step_optimizer(self.optimizer)
        return {"loss": losses["loss_box_reg"].detach(), "loss_classifier": losses["loss_classifier"].detach(),
                "loss_objectness": losses["loss_objectness"].detach()}

    def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
        images, targets = batch
        outputs = self.model(list(images), copy.deepcopy(list(targets)))
        return {"val_avg_iou": average_best_iou(outputs, list(targets))}
