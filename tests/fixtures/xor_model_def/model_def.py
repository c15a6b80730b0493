"""Self-contained model definition used by the end-to-end master tests (submitted like a user's directory)."""
import os

import numpy as np
import torch
import torch.nn as nn

from determined.pytorch import DataLoader, PyTorchTrial, PyTorchTrialContext


class XorData(torch.utils.data.Dataset):
    def __init__(self, n, seed):
        rng = np.random.RandomState(seed)
        self.x = rng.randint(0, 2, size=(n, 2)).astype(np.float32)
        self.y = (self.x[:, 0] != self.x[:, 1]).astype(np.float32).reshape(-1, 1)

    def __len__(self):
        return len(self.x)

    def __getitem__(self, i):
        return torch.from_numpy(self.x[i]), torch.from_numpy(self.y[i])


class XorTrial(PyTorchTrial):
    def __init__(self, context: PyTorchTrialContext) -> None:
        self.context = context
        h = int(context.get_hparam("hidden_size"))
        self.model = context.wrap_model(nn.Sequential(nn.Linear(2, h), nn.Sigmoid(), nn.Linear(h, 1), nn.Sigmoid()))
        self.opt = context.wrap_optimizer(torch.optim.SGD(self.model.parameters(), lr=float(context.get_hparam("learning_rate"))))
        self.loss_fn = nn.BCELoss()
        chaos = float(context.get_hparams().get("chaos_probability", 0.0))
        self.chaos = chaos
        if context.get_hparams().get("invalid_hp"):
            import determined as det

            raise det.InvalidHP("this hyperparameter point is declared invalid")

    def train_batch(self, batch, epoch_idx, batch_idx):
        if self.chaos and np.random.RandomState(os.getpid() + batch_idx).rand() < self.chaos:
            raise RuntimeError("chaos monkey strikes")
        x, y = batch
        loss = self.loss_fn(self.model(x), y)
        self.context.backward(loss)
        self.context.step_optimizer(self.opt)
        return {"loss": loss}

    def evaluate_batch(self, batch):
        x, y = batch
        out = self.model(x)
        return {"validation_loss": self.loss_fn(out, y), "error": ((out > 0.5).float() != y).float().mean()}

    def build_training_data_loader(self):
        return DataLoader(XorData(256, 1), batch_size=self.context.get_per_slot_batch_size(), shuffle=True)

    def build_validation_data_loader(self):
        return DataLoader(XorData(64, 2), batch_size=self.context.get_per_slot_batch_size())
