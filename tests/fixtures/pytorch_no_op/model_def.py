"""A PyTorchTrial that learns nothing and reports one draw from every random number generator the harness checkpoints
(numpy, ``random``, torch CPU) as validation metrics — the fixture behind the RNG-restore end-to-end test (role of the
reference's ``e2e_tests/tests/fixtures/pytorch_no_op``)."""
import random

import numpy as np
import torch
import torch.nn as nn

from determined.pytorch import DataLoader, PyTorchTrial, PyTorchTrialContext


class OnesDataset(torch.utils.data.Dataset):
    def __len__(self):
        return 64

    def __getitem__(self, i):
        return torch.ones(1), torch.ones(1)


class NoopPyTorchTrial(PyTorchTrial):
    def __init__(self, context: PyTorchTrialContext) -> None:
        self.context = context
        self.model = context.wrap_model(nn.Linear(1, 1, bias=False))
        self.opt = context.wrap_optimizer(torch.optim.SGD(self.model.parameters(), lr=0.001))

    def train_batch(self, batch, epoch_idx, batch_idx):
        x, y = batch
        loss = ((self.model(x) - y) ** 2).mean()
        self.context.backward(loss)
        self.context.step_optimizer(self.opt)
        return {"loss": loss}

    def evaluate_full_dataset(self, data_loader):
        return {"validation_loss": float(np.random.rand()), "np_rand": float(np.random.rand()), "rand_rand": random.random(),
                "torch_rand": float(torch.rand(1).item())}

    def build_training_data_loader(self):
        return DataLoader(OnesDataset(), batch_size=self.context.get_per_slot_batch_size())

    def build_validation_data_loader(self):
        return DataLoader(OnesDataset(), batch_size=self.context.get_per_slot_batch_size())
