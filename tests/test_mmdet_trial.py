"""mmdetection wrapper (reference: examples/computer_vision/mmdetection_pytorch): the parts that do not need mmdetection,
the clear failure without it, and the trial end to end against stand-in ``mmcv`` / ``mmdet`` modules that implement the
handful of entry points the wrapper calls."""
import sys
import types

import numpy as np
import pytest
import torch

from determined_b200.models import mmdet_trial as M
from tests.harness_utils import TrainAndValidate, make_controller


def test_warmup_wrapper_values():
    w = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([w], lr=0.1)
    sched = M.warmup_wrapper(torch.optim.lr_scheduler.MultiStepLR)("linear", 4, 0.001, opt, [6, 8], 0.1)
    lrs = []
    for _ in range(10):
        lrs.append(opt.param_groups[0]["lr"])
        opt.step()
        sched.step()
    want = [0.1 * (0.001 + k / 4 * 0.999) for k in (1, 2, 3, 4)]
    assert np.allclose(lrs[:4], want) and np.allclose(lrs[4:6], 0.1) and np.allclose(lrs[6:8], 0.01) and np.allclose(lrs[8:], 0.001)
    opt2 = torch.optim.SGD([w], lr=1.0)
    s2 = M.warmup_wrapper(torch.optim.lr_scheduler.MultiStepLR)("exp", 2, 0.25, opt2, [100], 0.1)
    assert np.isclose(opt2.param_groups[0]["lr"], 0.25 ** 0.5)
    s3 = M.warmup_wrapper(torch.optim.lr_scheduler.MultiStepLR)(None, 0, 1.0, torch.optim.SGD([w], lr=0.5), [1], 0.1)
    assert s3.get_last_lr() == [0.5]
    with pytest.raises(ValueError):
        M.warmup_wrapper(torch.optim.lr_scheduler.MultiStepLR)("cosine", 2, 0.5, opt2, [1], 0.1)
    del s2


def test_parse_losses():
    total, log = M.parse_losses({"loss_cls": torch.tensor([1.0, 3.0]), "loss_bbox": [torch.tensor(1.0), torch.tensor([2.0, 4.0])],
                                 "acc": torch.tensor(90.0)})
    assert total.item() == 2.0 + 1.0 + 3.0 and log["acc"].item() == 90.0 and set(log) == {"loss_cls", "loss_bbox", "acc"}
    with pytest.raises(TypeError):
        M.parse_losses({"loss": 1.0})


def test_group_batch_sampler_keeps_groups_together():
    flags = np.array([0] * 7 + [1] * 10)
    s = M.GroupBatchSampler(flags, samples_per_gpu=2, num_replicas=2, seed=1)
    batches = list(s)
    assert len(batches) == len(s) == 2 * 2 + 3 * 2                   # 7 → 8 and 10 → 12 samples, in batches of 2
    assert all(len(set(flags[b])) == 1 for b in batches)             # one aspect-ratio group per batch
    assert set(i for b in batches for i in b) == set(range(17))      # everything is visited
    for i in range(0, len(batches), 2):                              # the two ranks of a step draw from the same group
        assert flags[batches[i][0]] == flags[batches[i + 1][0]]
    assert list(s) != batches                                        # reshuffled every epoch


def test_trial_fails_clearly_without_mmdet():
    if "mmdet" in sys.modules:
        pytest.skip("mmdet is installed")
    with pytest.raises(ImportError, match="mmdet"):
        M._require_mmdet()


# ---- stand-ins ----------------------------------------------------------------------------------------------------------------
class _Cfg(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    @staticmethod
    def wrap(x):
        return _Cfg({k: _Cfg.wrap(v) for k, v in x.items()}) if isinstance(x, dict) else x


class _Container:
    def __init__(self, data):
        self.data = data


class _Detector(torch.nn.Module):
    def __init__(self, width):
        super().__init__()
        self.net = torch.nn.Linear(width, 1)

    def forward_train(self, img, gt):
        err = (self.net(img).squeeze(-1) - gt) ** 2
        return {"loss_reg": err, "loss_aux": [0.1 * err.mean()], "acc": torch.tensor(50.0)}

    def forward(self, return_loss=True, rescale=False, img=None, **kw):
        return [float(v) for v in self.net(img).squeeze(-1)]


class _Dataset(torch.utils.data.Dataset):
    CLASSES = ("thing",)

    def __init__(self, n, width):
        g = torch.Generator().manual_seed(n)
        self.x = torch.randn(n, width, generator=g)
        self.y = self.x @ torch.arange(1, width + 1, dtype=torch.float32) / width
        self.flag = (np.arange(n) % 3 == 0).astype(np.int64)
        self.evaluated = None

    def __len__(self):
        return len(self.x)

    def __getitem__(self, i):
        return {"img": self.x[i], "gt": self.y[i]}

    def evaluate(self, results, **kwargs):
        self.evaluated = (len(results), kwargs)
        err = float(np.mean((np.array(results) - self.y.numpy()) ** 2))
        return {"bbox_mAP": 1.0 / (1.0 + err)}


@pytest.fixture
def fake_mmdet(tmp_path, monkeypatch):
    cfg_file = tmp_path / "cfg.py"
    cfg_file.write_text("model = dict(width=6)\noptimizer = dict(lr=0.05, momentum=0.9)\ntrain_cfg = None\ntest_cfg = None\n"
                        "data = dict(train=dict(n=64, ann_file='x'), val=dict(n=16, ann_file='y'))\n"
                        "evaluation = dict(interval=1, metric='bbox')\n")

    def fromfile(path):
        ns = {}
        exec(open(path).read(), ns)
        return _Cfg.wrap({k: v for k, v in ns.items() if not k.startswith("__")})

    mmcv = types.ModuleType("mmcv")
    mmcv.Config = types.SimpleNamespace(fromfile=fromfile)
    runner = types.ModuleType("mmcv.runner")
    runner.build_optimizer = lambda model, c: torch.optim.SGD(model.parameters(), lr=c.lr, momentum=c.momentum)
    parallel = types.ModuleType("mmcv.parallel")
    parallel.collate = lambda batch, samples_per_gpu=1: {k: _Container([torch.stack([b[k] for b in batch])]) for k in batch[0]}
    mmdet = types.ModuleType("mmdet")
    models = types.ModuleType("mmdet.models")
    models.build_detector = lambda c, train_cfg=None, test_cfg=None: _Detector(c.width)
    datasets = types.ModuleType("mmdet.datasets")
    built = []

    def build_dataset(c):
        built.append(_Dataset(c.n, 6))
        return built[-1]

    datasets.build_dataset = build_dataset
    for name, mod in {"mmcv": mmcv, "mmcv.runner": runner, "mmcv.parallel": parallel, "mmdet": mmdet, "mmdet.models": models,
                      "mmdet.datasets": datasets}.items():
        monkeypatch.setitem(sys.modules, name, mod)
    return {"config_file": str(cfg_file), "built": built}


def test_trial_against_stand_in_mmdet(fake_mmdet):
    hp = {"global_batch_size": 8, "config_file": fake_mmdet["config_file"], "clip_grads": True, "clip_grads_norm": 5.0,
          "step1": 60, "step2": 70, "warmup": "linear", "warmup_iters": 5, "warmup_ratio": 0.01, "gamma": 0.1}
    tv = TrainAndValidate()
    controller = make_controller(M.MMDetTrial, hp, tv.send(steps=4, validation_freq=4, scheduling_unit=20),
                                 exp_config={"data": {"train_ann_file": "/ann/train.json", "val_ann_file": "/ann/val.json",
                                                      "workers_per_gpu": 0, "backend": "local"}})
    trial = controller.trial
    assert trial.cfg.data.train.ann_file == "/ann/train.json" and trial.cfg.data.val.test_mode is True
    controller.run()
    losses = [m["loss"] for m in tv.training_metrics]
    assert len(losses) == 80 and np.mean(losses[-10:]) < 0.2 * np.mean(losses[:10])
    assert {"loss_reg", "loss_aux", "acc"} <= set(tv.training_metrics[0])
    assert np.isclose(trial.optimizer.param_groups[0]["lr"], 0.05 * 0.1 * 0.1)         # both milestones passed
    val = tv.validation_metrics[-1]
    assert 0.5 < val["bbox_mAP"] <= 1.0
    val_ds = fake_mmdet["built"][-1]
    assert val_ds.evaluated == (16, {"metric": "bbox"})                                # bookkeeping keys were dropped
    assert trial.model.CLASSES == ("thing",)
