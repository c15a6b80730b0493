"""``DETRTrial`` — DETR inside ``PyTorchTrial``.

Counterpart of the reference's ``examples/computer_vision/detr_coco_pytorch/model_def.py:36-219``: same hyperparameter
names, AdamW with a separate (lower) learning rate for the backbone, ``StepLR(lr_drop)`` stepped every epoch, gradient
clipping at ``clip_max_norm``, the weighted sum of the matched-set losses as the training loss with every component (and
its scaled value) reported as a metric, and ``evaluate_full_dataset`` producing the losses plus COCO box AP
(``mAP, mAP_50, mAP_75, mAP_small, mAP_medium, mAP_large``). The reference offers ``backend: fake`` configs
(``const_fake.yaml``) for machines without the COCO archive; that is the back-end provided here: synthetic scenes of
coloured rectangles whose colour encodes the class, in varying image sizes so the padding mask is exercised.
"""
from __future__ import annotations

from typing import Any, Dict, List, Tuple

import torch

from .. import pytorch as dp
from .detr import (DETR, HungarianMatcher, SetCriterion, box_xyxy_to_cxcywh, build_weight_dict, coco_average_precision,
                   post_process)


class FakeCocoDetection(torch.utils.data.Dataset):
    """``(image [3, h, w], {"boxes": cxcywh in [0, 1], "labels", "orig_size": (h, w), "image_id"})``."""

    def __init__(self, size: int, image_size: int, num_classes: int, seed: int, max_objects: int = 4) -> None:
        self.size, self.hw, self.num_classes, self.seed, self.max_objects = size, image_size, num_classes, seed, max_objects

    def __len__(self) -> int:
        return self.size

    def __getitem__(self, i: int) -> Tuple[torch.Tensor, Dict[str, torch.Tensor]]:
        g = torch.Generator().manual_seed(self.seed * 1_000_003 + i)
        r = lambda lo, hi: int(torch.randint(lo, hi, (1,), generator=g))      # noqa: E731
        h, w = r(self.hw * 3 // 4, self.hw + 1) // 32 * 32, r(self.hw * 3 // 4, self.hw + 1) // 32 * 32
        img = torch.rand(3, h, w, generator=g) * 0.2
        boxes, labels = [], []
        for _ in range(r(1, self.max_objects + 1)):
            bw, bh = r(w // 8, w // 2), r(h // 8, h // 2)
            x0, y0 = r(0, w - bw), r(0, h - bh)
            c = r(0, self.num_classes)
            hue = (c + 0.5) / self.num_classes * 6.283
            colour = 0.55 + 0.4 * torch.cos(torch.tensor([hue, hue + 2.094, hue + 4.189]))
            img[:, y0:y0 + bh, x0:x0 + bw] = colour[:, None, None]
            boxes.append([x0, y0, x0 + bw, y0 + bh])
            labels.append(c)
        b = box_xyxy_to_cxcywh(torch.tensor(boxes, dtype=torch.float32)) / torch.tensor([w, h, w, h], dtype=torch.float32)
        return img, {"boxes": b, "labels": torch.tensor(labels, dtype=torch.int64), "orig_size": torch.tensor([h, w]),
                     "image_id": torch.tensor(i)}


def pad_collate(batch: List[Tuple[torch.Tensor, Dict[str, torch.Tensor]]]):
    """Images of different sizes → one zero-padded tensor plus the padding mask (True where there is no image)."""
    H = max(im.shape[1] for im, _ in batch)
    W = max(im.shape[2] for im, _ in batch)
    images = torch.zeros(len(batch), 3, H, W)
    mask = torch.ones(len(batch), H, W, dtype=torch.bool)
    for i, (im, _) in enumerate(batch):
        images[i, :, :im.shape[1], :im.shape[2]] = im
        mask[i, :im.shape[1], :im.shape[2]] = False
    return images, mask, [t for _, t in batch]


class DETRTrial(dp.PyTorchTrial):
    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = dict(context.get_hparams())
        self.hp = hp
        if hp.get("backend", "fake") != "fake":
            raise ValueError("only `backend: fake` is available: the COCO archive cannot be downloaded here")
        self.num_classes = int(hp.get("num_classes", 91))
        hidden = int(hp.get("hidden_dim", 256))
        dec_layers, aux = int(hp.get("dec_layers", 6)), bool(hp.get("aux_loss", True))
        model = DETR(self.num_classes, int(hp.get("num_queries", 100)), hidden, int(hp.get("nheads", hidden // 64)),
                     int(hp.get("enc_layers", 6)), dec_layers, int(hp.get("dim_feedforward", 2048)), float(hp.get("dropout", 0.1)),
                     bool(hp.get("pre_norm", False)), str(hp.get("backbone", "resnet50")), aux)
        self.model = context.wrap_model(model)
        matcher = HungarianMatcher(float(hp.get("set_cost_class", 1)), float(hp.get("set_cost_bbox", 5)), float(hp.get("set_cost_giou", 2)))
        self.weight_dict = build_weight_dict(float(hp.get("bbox_loss_coef", 5)), float(hp.get("giou_loss_coef", 2)), dec_layers, aux)
        self.criterion = SetCriterion(self.num_classes, matcher, self.weight_dict, float(hp.get("eos_coef", 0.1)),
                                      context.distributed.get_size()).to(context.device)
        named = [(n, p) for n, p in self.model.named_parameters() if p.requires_grad]
        groups = [{"params": [p for n, p in named if "backbone" not in n]},
                  {"params": [p for n, p in named if "backbone" in n], "lr": float(hp.get("lr_backbone", 1e-5))}]
        self.optimizer = context.wrap_optimizer(torch.optim.AdamW(groups, lr=float(hp.get("lr", 1e-4)),
                                                                  weight_decay=float(hp.get("weight_decay", 1e-4))))
        self.lr_scheduler = context.wrap_lr_scheduler(torch.optim.lr_scheduler.StepLR(self.optimizer, int(hp.get("lr_drop", 100))),
                                                      dp.LRScheduler.StepMode.STEP_EVERY_EPOCH)
        clip = float(hp.get("clip_max_norm", 0.1))
        self.clip_grads = dp.ClipGradsL2Norm(clip) if clip > 0 else None

    def _loader(self, size: int, seed: int, shuffle: bool) -> dp.DataLoader:
        ds = FakeCocoDetection(size, int(self.hp.get("image_size", 256)), self.num_classes, seed, int(self.hp.get("max_objects", 4)))
        return dp.DataLoader(ds, batch_size=self.context.get_per_slot_batch_size(), collate_fn=pad_collate, shuffle=shuffle,
                             num_workers=int(self.hp.get("num_workers", 0)))

    def build_training_data_loader(self) -> dp.DataLoader:
        return self._loader(int(self.hp.get("dataset_size", 512)), 1, True)

    def build_validation_data_loader(self) -> dp.DataLoader:
        return self._loader(int(self.hp.get("validation_size", 64)), 2, False)

    def _loss_metrics(self, loss_dict: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        wd = self.weight_dict
        scaled = {f"{k}_scaled": v * wd[k] for k, v in loss_dict.items() if k in wd}
        out = dict(loss_dict)
        out["sum_unscaled"] = sum(loss_dict.values())
        out["sum_scaled"] = sum(scaled.values())
        out.update(scaled)
        return out

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        images, mask, targets = batch
        loss_dict = self.criterion(self.model(images, mask), targets)
        loss = sum(loss_dict[k] * self.weight_dict[k] for k in loss_dict if k in self.weight_dict)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer, clip_grads=self.clip_grads)
        out = self._loss_metrics(loss_dict)
        out["loss"] = loss
        return {k: v.detach() for k, v in out.items()}

    def evaluate_full_dataset(self, data_loader: torch.utils.data.DataLoader) -> Dict[str, Any]:
        totals: Dict[str, float] = {}
        detections, truth = {}, {}
        n = 0
        with torch.no_grad():
            for batch in data_loader:
                images, mask, targets = self.context.to_device(batch)
                with self.context.autocast():
                    outputs = self.model(images, mask)
                for k, v in self._loss_metrics(self.criterion(outputs, targets)).items():
                    totals[k] = totals.get(k, 0.0) + float(v)
                n += 1
                sizes = torch.stack([t["orig_size"] for t in targets]).float()
                for t, res, (h, w) in zip(targets, post_process(outputs, sizes), sizes.tolist()):
                    iid = int(t["image_id"])
                    detections[iid] = {k: v.float().cpu().numpy() for k, v in res.items()}
                    cx, cy, bw, bh = (t["boxes"].float().cpu() * torch.tensor([w, h, w, h])).unbind(-1)
                    truth[iid] = {"boxes": torch.stack([cx - bw / 2, cy - bh / 2, cx + bw / 2, cy + bh / 2], -1).numpy(),
                                  "labels": t["labels"].cpu().numpy()}
        metrics: Dict[str, Any] = {k: v / max(n, 1) for k, v in totals.items()}
        metrics.update(coco_average_precision(detections, truth))
        return metrics
