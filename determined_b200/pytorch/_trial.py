"""``PyTorchTrial`` — the user-facing trial definition.

Parity: reference ``harness/determined/pytorch/_pytorch_trial.py:769-1020``. A trial builds its models,
optimizers and LR schedulers in ``__init__`` through the context's ``wrap_*`` calls, then implements
``train_batch`` and either ``evaluate_batch`` or ``evaluate_full_dataset``.
"""
from __future__ import annotations

import abc
from typing import Any, Dict, Union

import torch

from .._trial import Trial
from . import _callback, _data, _reducer
from ._context import PyTorchTrialContext


class PyTorchTrial(Trial):
    trial_context_class = PyTorchTrialContext

    @abc.abstractmethod
    def __init__(self, context: PyTorchTrialContext) -> None:
        """Wrap models, optimizers and LR schedulers here::

            self.context = context
            self.model = context.wrap_model(MyModel())
            self.opt = context.wrap_optimizer(torch.optim.AdamW(self.model.parameters(), lr=...))
        """

    @abc.abstractmethod
    def train_batch(self, batch: _data.TorchData, epoch_idx: int, batch_idx: int) -> Union[torch.Tensor, Dict[str, Any]]:
        """Forward, ``context.backward(loss)``, ``context.step_optimizer(opt)``; return the loss tensor or a
        dict of metric name → scalar tensor (must contain ``"loss"`` when a dict)."""

    @abc.abstractmethod
    def build_training_data_loader(self) -> _data.DataLoader:
        ...

    @abc.abstractmethod
    def build_validation_data_loader(self) -> _data.DataLoader:
        ...

    def build_callbacks(self) -> Dict[str, _callback.PyTorchCallback]:
        return {}

    def evaluate_batch(self, batch: _data.TorchData) -> Dict[str, Any]:
        """Per-batch validation metrics; reduced by ``evaluation_reducer()`` across batches and ranks."""
        raise NotImplementedError

    def evaluation_reducer(self) -> Union[_reducer.Reducer, _reducer.MetricReducer,
                                          Dict[str, Union[_reducer.Reducer, _reducer.MetricReducer]]]:
        return _reducer.Reducer.AVG

    def evaluate_full_dataset(self, data_loader: torch.utils.data.DataLoader) -> Dict[str, Any]:
        """Whole-dataset validation on the chief only (mutually exclusive with ``evaluate_batch``)."""
        raise NotImplementedError

    # ---- deprecated single-model API of earlier releases, still honoured when overridden ----
    def build_model(self) -> torch.nn.Module:
        raise NotImplementedError

    def optimizer(self, model: torch.nn.Module) -> torch.optim.Optimizer:
        raise NotImplementedError

    def create_lr_scheduler(self, optimizer: torch.optim.Optimizer):
        return None


def reset_parameters(model: torch.nn.Module) -> None:
    """Re-initialise every child module that defines ``reset_parameters`` (reference: ``:1029-1042``)."""
    missing = [type(m).__name__ for m in model.children() if not hasattr(m, "reset_parameters")]
    if missing:
        raise AttributeError(f"reset_parameters() missing on: {sorted(set(missing))}")
    for m in model.children():
        m.reset_parameters()  # type: ignore[operator]
