"""``MMDetTrial`` — any detector of the mmdetection model zoo inside ``PyTorchTrial`` (gated on the ``mmdet`` / ``mmcv``
packages, which this image does not ship).

Counterpart of the reference's ``examples/computer_vision/mmdetection_pytorch`` (``model_def.py:28-141``,
``utils/lr_schedulers.py:6-38``, ``utils/data.py``): the experiment names an mmdetection config file
(``hyperparameters.config_file``); the trial loads it, points the train / validation annotation files at
``data.train_ann_file`` / ``data.val_ann_file``, builds the detector and the optimizer the config describes, trains with
``forward_train`` + the summed ``loss*`` entries, steps a ``MultiStepLR`` with warm-up every batch, and evaluates the whole
validation set with the dataset's own ``evaluate`` (COCO ``bbox_mAP``). Like any stock ``torch.nn.Module`` (see
``detection_trial.py``) the detector goes through ``wrap_model`` / ``wrap_optimizer`` unchanged: flat-bucket fused
optimizer, in-kernel gradient all-reduce, checkpoints and searchers apply; the network itself runs on torch's kernels.

The pieces that do not need mmdetection — the warm-up scheduler wrapper, the loss parser, the group-by-aspect-ratio batch
sampler — are plain functions here and are tested without it; the trial raises a clear error when the packages are absent.
"""
from __future__ import annotations

import math
from typing import Any, Dict, Iterator, List, Sequence, Tuple, Type

import numpy as np
import torch

from .. import pytorch as dp


def warmup_wrapper(scheduler_type: Type[torch.optim.lr_scheduler.LRScheduler]) -> Type[torch.optim.lr_scheduler.LRScheduler]:
    """``scheduler_type`` with a warm-up phase in front: for the first ``warmup_iters`` steps the learning rate is the base
    rate times ``warmup_ratio`` (``constant``), a line from ``warmup_ratio`` to 1 (``linear``) or ``warmup_ratio ** (1 -
    progress)`` (``exp``); afterwards the wrapped schedule applies. Constructor: ``(warmup, warmup_iters, warmup_ratio,
    optimizer, *scheduler_args)``."""

    class Warmup(scheduler_type):  # type: ignore[misc, valid-type]
        def __init__(self, warmup, warmup_iters, warmup_ratio, *args, **kwargs):
            if warmup is not None:
                if warmup not in ("constant", "linear", "exp"):
                    raise ValueError("warmup must be constant, linear, exp or None")
                if not warmup_iters > 0 or not 0 < warmup_ratio <= 1.0:
                    raise ValueError("warmup needs warmup_iters > 0 and 0 < warmup_ratio <= 1")
            self.warmup, self.warmup_iters, self.warmup_ratio = warmup, int(warmup_iters or 0), float(warmup_ratio or 1.0)
            super().__init__(*args, **kwargs)

        def warmup_multiplier(self) -> float:
            progress = (self.last_epoch + 1) / self.warmup_iters
            if self.warmup == "constant":
                return self.warmup_ratio
            if self.warmup == "linear":
                return self.warmup_ratio + progress * (1 - self.warmup_ratio)
            return self.warmup_ratio ** (1 - progress)

        def get_lr(self):
            if self.warmup is not None and self.last_epoch < self.warmup_iters:
                m = self.warmup_multiplier()
                return [m * b for b in self.base_lrs]
            return super().get_lr()

    Warmup.__name__ = f"Warmup{scheduler_type.__name__}"
    return Warmup


def parse_losses(losses: Dict[str, Any]) -> Tuple[torch.Tensor, Dict[str, torch.Tensor]]:
    """mmdetection's loss dictionary → (total, log values): every entry is averaged (lists of tensors are summed over
    their means); the total is the sum of the entries whose key contains ``loss``."""
    log: Dict[str, torch.Tensor] = {}
    for name, value in losses.items():
        if isinstance(value, torch.Tensor):
            log[name] = value.mean()
        elif isinstance(value, (list, tuple)):
            log[name] = sum(v.mean() for v in value)
        else:
            raise TypeError(f"{name} is not a tensor or a list of tensors")
    total = sum(v for k, v in log.items() if "loss" in k)
    return total, {k: v.detach() for k, v in log.items()}


class GroupBatchSampler(torch.utils.data.Sampler):
    """Batches of ``samples_per_gpu`` images that share an aspect-ratio group (``flags[i]`` is 0 for portrait, 1 for
    landscape, as mmdetection datasets report), so that padding inside a batch stays small. Each group is shuffled, padded
    up to a multiple of ``samples_per_gpu * num_replicas`` by repeating some of its members, cut into batches, and the
    batches of all groups are shuffled together; the framework's ``DataLoader`` then hands every ``num_replicas``-th batch
    to each rank."""

    def __init__(self, flags: Sequence[int], samples_per_gpu: int, num_replicas: int = 1, seed: int = 0) -> None:
        self.flags, self.n, self.replicas = np.asarray(flags), int(samples_per_gpu), int(num_replicas)
        self.seed, self.epoch = seed, 0
        chunk = self.n * self.replicas
        self.num_batches = sum(int(math.ceil(c / chunk)) * self.replicas for c in np.bincount(self.flags) if c)

    def __len__(self) -> int:
        return self.num_batches

    def __iter__(self) -> Iterator[List[int]]:
        rng = np.random.RandomState(self.seed + self.epoch)
        self.epoch += 1
        chunk = self.n * self.replicas
        batches: List[List[int]] = []
        for g in np.unique(self.flags):
            idx = np.where(self.flags == g)[0]
            rng.shuffle(idx)
            extra = int(math.ceil(len(idx) / chunk)) * chunk - len(idx)
            idx = np.concatenate([idx, rng.choice(idx, extra)]) if extra else idx
            batches += [idx[i:i + self.n].tolist() for i in range(0, len(idx), self.n)]
        # shuffle whole groups of `replicas` consecutive batches, so the ranks of one step draw from the same group
        steps = [batches[i:i + self.replicas] for i in range(0, len(batches), self.replicas)]
        for j in rng.permutation(len(steps)):
            yield from steps[j]


def unwrap_containers(batch: Dict[str, Any]) -> Dict[str, Any]:
    """mmcv's collate wraps every field in a ``DataContainer`` holding one list per device; take this device's, and put
    the image tensor first so that the batch length can be read off it."""
    out = {k: (v.data[0] if hasattr(v, "data") and not isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
    if "img" in out:
        out = dict([("img", out.pop("img"))], **out)
    return out


def _require_mmdet():
    try:
        import mmcv  # noqa: F401
        import mmdet  # noqa: F401
    except ImportError as e:
        raise ImportError("MMDetTrial needs the `mmdet` and `mmcv` packages (pip install mmdet mmcv-full in the task "
                          "environment); they are not part of this image") from e


class MMDetTrial(dp.PyTorchTrial):
    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        _require_mmdet()
        from mmcv import Config
        from mmcv.runner import build_optimizer
        from mmdet.models import build_detector

        self.context = context
        self.hparams = dict(context.get_hparams())
        self.data_config = dict(context.get_data_config() or {})
        cfg = Config.fromfile(self.hparams["config_file"])
        cfg.data.train.ann_file = self.data_config.get("train_ann_file", cfg.data.train.get("ann_file"))
        cfg.data.val.ann_file = self.data_config.get("val_ann_file", cfg.data.val.get("ann_file"))
        cfg.data.val.test_mode = True
        cfg.data.workers_per_gpu = int(self.data_config.get("workers_per_gpu", 2))
        self.cfg = cfg
        self.model = context.wrap_model(build_detector(cfg.model, train_cfg=cfg.get("train_cfg"), test_cfg=cfg.get("test_cfg")))
        self.optimizer = context.wrap_optimizer(build_optimizer(self.model, cfg.optimizer))
        sched_cls = warmup_wrapper(torch.optim.lr_scheduler.MultiStepLR)
        hp = self.hparams
        self.scheduler = context.wrap_lr_scheduler(
            sched_cls(hp.get("warmup"), hp.get("warmup_iters", 0), hp.get("warmup_ratio", 1.0), self.optimizer,
                      [int(hp["step1"]), int(hp["step2"])], float(hp.get("gamma", 0.1))),
            step_mode=dp.LRScheduler.StepMode.MANUAL_STEP)
        self.clip_grads = dp.ClipGradsL2Norm(float(hp["clip_grads_norm"])) if hp.get("clip_grads") else None

    def _loader(self, data_cfg: Any, samples_per_gpu: int, replicas: int, workers: int, train: bool) -> dp.DataLoader:
        from mmcv.parallel import collate
        from mmdet.datasets import build_dataset

        ds = build_dataset(data_cfg)
        self.model.CLASSES = getattr(ds, "CLASSES", None)
        flags = getattr(ds, "flag", np.zeros(len(ds), dtype=np.int64))
        if train:
            sampler: Any = GroupBatchSampler(flags, samples_per_gpu, replicas, seed=self.context.get_trial_seed())
        else:
            sampler = torch.utils.data.BatchSampler(torch.utils.data.SequentialSampler(ds), samples_per_gpu, drop_last=False)
        return dp.DataLoader(ds, batch_sampler=sampler, num_workers=workers,
                             collate_fn=lambda b: unwrap_containers(collate(b, samples_per_gpu=samples_per_gpu)))

    def build_training_data_loader(self) -> dp.DataLoader:
        return self._loader(self.cfg.data.train, self.context.get_per_slot_batch_size(), self.context.distributed.get_size(),
                            self.cfg.data.workers_per_gpu, True)

    def build_validation_data_loader(self) -> dp.DataLoader:
        return self._loader(self.cfg.data.val, 1, 1, self.cfg.data.workers_per_gpu, False)

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        loss, log_vars = parse_losses(self.model.forward_train(**batch))
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer, clip_grads=self.clip_grads)
        self.scheduler.step()
        return dict(log_vars, loss=loss.detach())

    def evaluate_full_dataset(self, data_loader: torch.utils.data.DataLoader) -> Dict[str, Any]:
        if self.data_config.get("backend") == "fake":
            return {"bbox_mAP": 0.0}
        results: List[Any] = []
        for batch in data_loader:
            batch = self.context.to_device(batch)
            with torch.no_grad():
                result = self.model(return_loss=False, rescale=True, **batch)
            if result and isinstance(result[0], tuple):
                from mmdet.core import encode_mask_results

                result = [(boxes, encode_mask_results(masks)) for boxes, masks in result]
            results.extend(result)
        kwargs = dict(self.cfg.get("evaluation", {}))
        for key in ("interval", "tmpdir", "start", "gpu_collect"):
            kwargs.pop(key, None)
        return dict(data_loader.dataset.evaluate(results, **kwargs))
