"""DETR (reference: examples/computer_vision/detr_coco_pytorch): box utilities, the Hungarian matcher, the set criterion,
COCO-style AP, and the trial end to end on the synthetic back-end."""
import numpy as np
import torch

from determined_b200.models import detr
from determined_b200.models.detr_trial import DETRTrial, FakeCocoDetection, pad_collate
from tests.harness_utils import TrainAndValidate, make_controller

HP = {"global_batch_size": 4, "lr": 2e-3, "lr_backbone": 2e-4, "weight_decay": 1e-4, "lr_drop": 100, "clip_max_norm": 0.1,
      "backbone": "resnet18", "enc_layers": 1, "dec_layers": 2, "dim_feedforward": 128, "hidden_dim": 64, "nheads": 1,
      "dropout": 0.0, "num_queries": 8, "aux_loss": True, "num_classes": 3, "image_size": 96, "dataset_size": 16,
      "validation_size": 4, "max_objects": 2, "backend": "fake"}


def test_box_ops_and_giou():
    a = torch.tensor([[0.0, 0.0, 2.0, 2.0]])
    b = torch.tensor([[1.0, 1.0, 3.0, 3.0], [4.0, 4.0, 5.0, 5.0], [0.0, 0.0, 2.0, 2.0]])
    iou, union = detr.box_iou(a, b)
    assert torch.allclose(iou, torch.tensor([[1 / 7, 0.0, 1.0]])) and torch.allclose(union, torch.tensor([[7.0, 5.0, 4.0]]))
    g = detr.generalized_box_iou(a, b)
    assert torch.allclose(g, torch.tensor([[1 / 7 - 2 / 9, -20 / 25, 1.0]]), atol=1e-6)
    c = torch.rand(5, 4) * 0.4 + 0.1
    assert torch.allclose(detr.box_xyxy_to_cxcywh(detr.box_cxcywh_to_xyxy(c)), c, atol=1e-6)


def test_sine_position_embedding_respects_padding():
    mask = torch.zeros(2, 4, 6, dtype=torch.bool)
    mask[1, :, 3:] = True                                 # second image is half as wide
    pos = detr.sine_position_embedding(mask, 16).view(2, 4, 6, 32)
    assert pos.shape == (2, 4, 6, 32) and pos.abs().max() <= 1.0
    # the last valid column of either image sits at the same normalised x (2π), so its x-embedding is identical
    assert torch.allclose(pos[0, 0, 5, 16:], pos[1, 0, 2, 16:], atol=1e-4)
    assert not torch.allclose(pos[0, 0, 2, 16:], pos[1, 0, 2, 16:], atol=1e-3)


def test_matcher_finds_the_permutation_and_criterion_is_zero_on_a_perfect_prediction():
    torch.manual_seed(0)
    tb = torch.tensor([[0.3, 0.3, 0.2, 0.2], [0.7, 0.6, 0.3, 0.2], [0.5, 0.5, 0.1, 0.4]])
    tl = torch.tensor([2, 0, 1])
    Q, C = 6, 3
    perm = torch.tensor([4, 1, 5])                        # query perm[i] predicts target i
    logits = torch.full((1, Q, C + 1), -20.0)
    logits[0, :, C] = 20.0
    boxes = torch.rand(1, Q, 4) * 0.1 + 0.05
    for i, q in enumerate(perm):
        logits[0, q] = -20.0
        logits[0, q, tl[i]] = 20.0
        boxes[0, q] = tb[i]
    out = {"pred_logits": logits, "pred_boxes": boxes}
    targets = [{"labels": tl, "boxes": tb}]
    (src, tgt), = detr.HungarianMatcher()(out, targets)
    assert dict(zip(tgt.tolist(), src.tolist())) == {0: 4, 1: 1, 2: 5}
    crit = detr.SetCriterion(C, detr.HungarianMatcher(), detr.build_weight_dict(5, 2, 1, False))
    losses = crit(out, targets)
    assert losses["loss_ce"] < 1e-6 and losses["loss_bbox"] < 1e-6 and losses["loss_giou"] < 1e-6
    assert losses["class_error"] == 0 and losses["cardinality_error"] == 0
    # an image without objects: only the classification term is non-trivial, nothing crashes
    empty = crit(out, [{"labels": torch.zeros(0, dtype=torch.int64), "boxes": torch.zeros(0, 4)}])
    assert torch.isfinite(empty["loss_ce"]) and empty["loss_bbox"] == 0


def test_weight_dict_covers_aux_layers():
    wd = detr.build_weight_dict(5, 2, 3, True)
    assert wd["loss_bbox_1"] == 5 and wd["loss_giou_0"] == 2 and wd["loss_ce_1"] == 1 and len(wd) == 9


def test_coco_average_precision_by_hand():
    gt = {0: {"boxes": np.array([[10, 10, 60, 60], [100, 100, 220, 220.0]]), "labels": np.array([1, 2])},
          1: {"boxes": np.array([[5, 5, 25, 25.0]]), "labels": np.array([1])}}
    perfect = {k: {"boxes": v["boxes"], "labels": v["labels"], "scores": np.ones(len(v["labels"]))} for k, v in gt.items()}
    assert set(detr.coco_average_precision(perfect, gt).values()) == {1.0}
    # a confident false positive of class 1 ahead of one of its two true positives: P/R = (1, .5) (.5, .5) (.67, 1)
    dets = {0: {"boxes": np.array([[300, 300, 340, 340], [10, 10, 60, 60], [100, 100, 220, 220.0]]), "labels": np.array([1, 1, 2]),
                "scores": np.array([0.9, 0.8, 0.7])}, 1: perfect[1]}
    ap = detr.coco_average_precision(dets, gt)
    want_c1 = (51 * 1.0 + 50 * (2 / 3)) / 101
    assert abs(ap["mAP"] - (want_c1 + 1.0) / 2) < 1e-9 and ap["mAP_large"] == 1.0 and abs(ap["mAP_medium"] - 0.5) < 1e-9
    assert detr.coco_average_precision({}, gt)["mAP"] == 0.0


def test_collate_pads_and_masks():
    ds = FakeCocoDetection(8, 96, 3, seed=0)
    images, mask, targets = pad_collate([ds[i] for i in range(4)])
    assert images.shape[0] == 4 and mask.shape == images.shape[:1] + images.shape[2:]
    for i, t in enumerate(targets):
        h, w = t["orig_size"].tolist()
        assert not mask[i, :h, :w].any() and mask[i].sum() == mask[i].numel() - h * w
        assert (t["boxes"] > 0).all() and (t["boxes"] < 1).all()


def test_detr_trial_trains_and_reports_coco_metrics():
    tv = TrainAndValidate()
    make_controller(DETRTrial, HP, tv.send(steps=3, validation_freq=3, scheduling_unit=8)).run()
    losses = [m["loss"] for m in tv.training_metrics]
    assert len(losses) == 24 and all(np.isfinite(losses))
    assert sum(losses[-6:]) < sum(losses[:6]), losses
    first = tv.training_metrics[0]
    assert {"loss_ce", "loss_bbox", "loss_giou", "loss_ce_0", "loss_bbox_scaled", "sum_scaled", "class_error"} <= set(first)
    assert abs(first["sum_scaled"] - first["loss"]) < 1e-3 * max(1.0, abs(first["loss"]))
    val = tv.validation_metrics[-1]
    assert {"mAP", "mAP_50", "mAP_75", "mAP_small", "mAP_medium", "mAP_large", "loss_ce", "sum_scaled"} <= set(val)
    assert 0.0 <= val["mAP_50"] <= 1.0


def test_frozen_backbone_parts_stay_frozen():
    m = detr.DETR(3, 4, 64, 1, 1, 1, 64, 0.0, False, "resnet18", False)
    m.train()
    assert all(not mod.training for mod in m.backbone.body.modules() if mod.__class__.__name__.startswith("BatchNorm"))
    trainable = [n for n, p in m.backbone.named_parameters() if p.requires_grad]
    assert trainable and all(n.startswith(("body.layer2", "body.layer3", "body.layer4")) and "bn" not in n for n in trainable)
