"""``PyTorchCallback`` hooks and the gradient-clipping helpers.

Parity: reference ``harness/determined/pytorch/_callback.py:7-137`` (hook names, ``state_dict`` /
``load_state_dict`` for checkpointing, ``ClipGradsL2Norm`` / ``ClipGradsL2Value``).

The clip helpers double as *markers*: ``PyTorchTrialContext.step_optimizer(opt, clip_grads=...)``
recognises them and routes L2-norm clipping through the fused grad-norm + update kernels instead of
``torch.nn.utils.clip_grad_norm_`` (one pass over the bucket, no host sync). Called directly they behave
like the reference's callbacks.
"""
from __future__ import annotations

import pathlib
from typing import Any, Dict, Iterable

import torch


class PyTorchCallback:
    """Override any subset. All hooks run on every rank unless noted."""

    def on_validation_start(self) -> None:
        pass

    def on_validation_end(self, metrics: Dict[str, Any]) -> None:
        """Runs on every rank; metrics are broadcast from the chief when this is overridden."""

    def on_checkpoint_end(self, checkpoint_dir: str) -> None:
        """Chief only, after ``state_dict.pth`` and ``code/`` are written."""

    def on_train_step_start(self, step_id: int) -> None:
        pass

    def on_train_step_end(self, step_id: int, metrics: Any) -> None:
        pass

    # deprecated hooks kept for API compatibility with existing model definitions
    def on_before_optimizer_step(self, parameters: Iterable[torch.Tensor]) -> None:
        pass

    def on_validation_step_start(self) -> None:
        pass

    def on_validation_step_end(self, metrics: Dict[str, Any]) -> None:
        pass

    def state_dict(self) -> Dict[str, Any]:
        return {}

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        pass


class ClipGradsL2Norm(PyTorchCallback):
    """Clip the global L2 norm of all gradients to ``clip_value``."""

    def __init__(self, clip_value: float) -> None:
        self.clip_value = float(clip_value)

    def on_before_optimizer_step(self, parameters: Iterable[torch.Tensor]) -> None:
        torch.nn.utils.clip_grad_norm_(list(parameters), self.clip_value)

    def __call__(self, parameters: Iterable[torch.Tensor]) -> None:
        self.on_before_optimizer_step(parameters)


class ClipGradsL2Value(PyTorchCallback):
    """Clamp every gradient element to ``[-clip_value, clip_value]``."""

    def __init__(self, clip_value: float) -> None:
        self.clip_value = float(clip_value)

    def on_before_optimizer_step(self, parameters: Iterable[torch.Tensor]) -> None:
        torch.nn.utils.clip_grad_value_(list(parameters), self.clip_value)

    def __call__(self, parameters: Iterable[torch.Tensor]) -> None:
        self.on_before_optimizer_step(parameters)


def is_overridden(full_method: Any, parent_class: type) -> bool:
    """True when the bound method comes from a subclass rather than ``parent_class`` itself."""
    if callable(full_method):
        return full_method.__func__ is not getattr(parent_class, full_method.__name__, None)  # type: ignore
    return False
