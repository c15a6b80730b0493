"""Tiny trials used by the harness tests (the same roles as the reference's
``harness/tests/experiment/fixtures/pytorch_xor_model.py`` and ``pytorch_onevar_model.py``; authored anew)."""
from typing import Any, Dict

import numpy as np
import torch
import torch.nn as nn

from determined_b200 import pytorch as dp


class _OneVarData(torch.utils.data.Dataset):
    """y = 2x on x = 0..n-1: with w initialised to 0 and SGD the weight after each batch is known in closed form."""

    def __init__(self, n: int = 64) -> None:
        self.n = n

    def __len__(self) -> int:
        return self.n

    def __getitem__(self, i: int):
        x = torch.tensor([float(i % 8)])
        return x, 2.0 * x


class OneVarTrial(dp.PyTorchTrial):
    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        model = nn.Linear(1, 1, bias=False)
        model.weight.data.fill_(0.0)
        self.model = context.wrap_model(model)
        self.lr = float(context.get_hparam("learning_rate"))
        self.opt = context.wrap_optimizer(torch.optim.SGD(self.model.parameters(), lr=self.lr))
        self.loss_fn = nn.MSELoss()

    def train_batch(self, batch, epoch_idx, batch_idx):
        x, y = batch
        w_before = self.model.weight.detach().clone().reshape(())
        loss = self.loss_fn(self.model(x), y)
        self.context.backward(loss)
        self.context.step_optimizer(self.opt)
        return {"loss": loss, "w_before": w_before, "w_after": self.model.weight.detach().clone().reshape(())}

    def evaluate_batch(self, batch):
        x, y = batch
        return {"val_loss": self.loss_fn(self.model(x), y)}

    def build_training_data_loader(self):
        return dp.DataLoader(_OneVarData(), batch_size=self.context.get_per_slot_batch_size())

    def build_validation_data_loader(self):
        return dp.DataLoader(_OneVarData(16), batch_size=self.context.get_per_slot_batch_size())


class _XorData(torch.utils.data.Dataset):
    def __init__(self, n: int = 64, seed: int = 0) -> None:
        rng = np.random.RandomState(seed)
        self.x = rng.randint(0, 2, size=(n, 2)).astype(np.float32)
        self.y = (self.x[:, 0] != self.x[:, 1]).astype(np.float32).reshape(-1, 1)

    def __len__(self) -> int:
        return len(self.x)

    def __getitem__(self, i: int):
        return {"data": torch.from_numpy(self.x[i]), "label": torch.from_numpy(self.y[i])}


class XorNet(nn.Module):
    def __init__(self, hidden: int = 8) -> None:
        super().__init__()
        self.l1 = nn.Linear(2, hidden)
        self.l2 = nn.Linear(hidden, 1)

    def forward(self, x):
        return torch.sigmoid(self.l2(torch.sigmoid(self.l1(x))))


class XorTrial(dp.PyTorchTrial):
    """Dict batches, dropout-free MLP, Adam, per-batch LR scheduler; metrics loss + accuracy."""

    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        self.model = context.wrap_model(XorNet(int(context.get_hparam("hidden_size"))))
        self.opt = context.wrap_optimizer(torch.optim.Adam(self.model.parameters(),
                                                           lr=float(context.get_hparam("learning_rate"))))
        self.sched = context.wrap_lr_scheduler(torch.optim.lr_scheduler.StepLR(self.opt, 5, 0.9),
                                               dp.LRScheduler.StepMode.STEP_EVERY_BATCH)
        self.loss_fn = nn.BCELoss()
        self.shuffle = bool(context.get_hparams().get("shuffle", True))

    def train_batch(self, batch, epoch_idx, batch_idx):
        out = self.model(batch["data"])
        loss = self.loss_fn(out, batch["label"])
        self.context.backward(loss)
        self.context.step_optimizer(self.opt)
        return {"loss": loss, "lr": torch.tensor(self.opt.param_groups[0]["lr"])}

    def evaluate_batch(self, batch):
        out = self.model(batch["data"])
        err = ((out > 0.5).float() != batch["label"]).float().mean()
        return {"loss": self.loss_fn(out, batch["label"]), "binary_error": err}

    def build_training_data_loader(self):
        return dp.DataLoader(_XorData(64, 1), batch_size=self.context.get_per_slot_batch_size(), shuffle=self.shuffle)

    def build_validation_data_loader(self):
        return dp.DataLoader(_XorData(32, 2), batch_size=self.context.get_per_slot_batch_size())


class XorTrialPerMetricReducers(XorTrial):
    def evaluation_reducer(self) -> Dict[str, Any]:
        return {"loss": dp.Reducer.AVG, "binary_error": dp.Reducer.MAX}


class _CountReducer(dp.MetricReducer):
    def __init__(self) -> None:
        self.total = 0.0
        self.n = 0

    def reset(self) -> None:
        self.total, self.n = 0.0, 0

    def accumulate(self, metric):
        self.total += float(metric)
        self.n += 1
        return metric

    def per_slot_reduce(self):
        return self.total, self.n

    def cross_slot_reduce(self, per_slot_metrics):
        t = sum(x[0] for x in per_slot_metrics)
        n = sum(x[1] for x in per_slot_metrics)
        return t / max(n, 1)


class XorTrialCustomReducer(XorTrial):
    def evaluation_reducer(self) -> Dict[str, Any]:
        return {"loss": _CountReducer(), "binary_error": dp.Reducer.AVG}


class XorTrialFullDataset(XorTrial):
    evaluate_batch = dp.PyTorchTrial.evaluate_batch  # not overridden

    def evaluate_full_dataset(self, data_loader):
        losses = []
        for batch in data_loader:
            batch = self.context.to_device(batch)
            losses.append(self.loss_fn(self.model(batch["data"]), batch["label"]))
        return {"loss": torch.stack(losses).mean(), "binary_error": 0.0}


class Counter(dp.PyTorchCallback):
    def __init__(self) -> None:
        self.validation_steps_started = 0
        self.validation_steps_ended = 0
        self.checkpoints_ended = 0

    def on_validation_start(self) -> None:
        self.validation_steps_started += 1

    def on_validation_end(self, metrics) -> None:
        self.validation_steps_ended += 1

    def on_checkpoint_end(self, checkpoint_dir: str) -> None:
        self.checkpoints_ended += 1

    def state_dict(self):
        return dict(self.__dict__)

    def load_state_dict(self, sd) -> None:
        self.__dict__.update(sd)


class XorTrialCallbacks(XorTrial):
    def __init__(self, context) -> None:
        super().__init__(context)
        self.counter = Counter()

    def build_callbacks(self):
        return {"counter": self.counter}


class XorTrialGradClipping(XorTrial):
    def train_batch(self, batch, epoch_idx, batch_idx):
        out = self.model(batch["data"])
        loss = self.loss_fn(out, batch["label"])
        self.context.backward(loss)
        clip = self.context.get_hparams().get("clip_norm")
        self.context.step_optimizer(self.opt, clip_grads=dp.ClipGradsL2Norm(clip) if clip else None)
        return {"loss": loss}


class LegacyXorTrial(dp.PyTorchTrial):
    """Deprecated single-model API: build_model / optimizer / train_batch(batch, model, epoch, idx)."""

    def __init__(self, context) -> None:
        self.context = context

    def build_model(self):
        return XorNet(8)

    def optimizer(self, model):
        return torch.optim.SGD(model.parameters(), lr=0.1)

    def train_batch(self, batch, model, epoch_idx, batch_idx):  # type: ignore[override]
        return {"loss": nn.functional.binary_cross_entropy(model(batch["data"]), batch["label"])}

    def evaluate_batch(self, batch, model):  # type: ignore[override]
        return {"loss": nn.functional.binary_cross_entropy(model(batch["data"]), batch["label"])}

    def build_training_data_loader(self):
        return dp.DataLoader(_XorData(64, 1), batch_size=self.context.get_per_slot_batch_size())

    def build_validation_data_loader(self):
        return dp.DataLoader(_XorData(32, 2), batch_size=self.context.get_per_slot_batch_size())


class XorTrialNonScalarValidation(XorTrial):
    """``evaluate_full_dataset`` that also returns a non-scalar (the per-example predictions), as the reference's
    ``XORTrialWithNonScalarValidation`` (``tests/experiment/fixtures/pytorch_xor_model.py:193-240``)."""

    evaluate_batch = None  # type: ignore[assignment]   exactly one of the two evaluation methods may be defined

    def evaluate_full_dataset(self, data_loader):
        preds, err, n = [], 0.0, 0
        for batch in data_loader:
            batch = self.context.to_device(batch)
            out = self.model(batch["data"])
            preds.append(out.detach().cpu())
            err += float(((out > 0.5).float() != batch["label"]).float().sum())
            n += out.shape[0]
        return {"binary_error": err / max(n, 1), "predictions": torch.cat(preds).numpy()}
