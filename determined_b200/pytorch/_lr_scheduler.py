"""``LRScheduler`` — a torch LR scheduler plus the policy of WHEN the harness steps it.

Parity: reference ``harness/determined/pytorch/_lr_scheduler.py:9-71``: ``STEP_EVERY_EPOCH`` steps after
the last batch of an epoch, ``STEP_EVERY_BATCH`` after every batch (every optimizer step when
``frequency`` > 1 is requested), ``MANUAL_STEP`` never.
"""
from __future__ import annotations

import enum
from typing import Any, Dict

import torch


class LRScheduler:
    class StepMode(enum.Enum):
        STEP_EVERY_EPOCH = 1
        STEP_EVERY_BATCH = 2
        MANUAL_STEP = 3

    def __init__(self, scheduler: Any, step_mode: "LRScheduler.StepMode", frequency: int = 1) -> None:
        if not hasattr(scheduler, "step"):
            raise TypeError("scheduler must expose .step()")
        if not isinstance(step_mode, LRScheduler.StepMode):
            raise TypeError("step_mode must be an LRScheduler.StepMode")
        if frequency < 1:
            raise ValueError("frequency must be >= 1")
        self._scheduler = scheduler
        self._step_mode = step_mode
        self._frequency = frequency

    def step(self, *args: Any, **kwargs: Any) -> None:
        self._scheduler.step(*args, **kwargs)

    def get_last_lr(self):
        return self._scheduler.get_last_lr()

    def state_dict(self) -> Dict[str, Any]:
        return self._scheduler.state_dict()

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        self._scheduler.load_state_dict(state_dict)

    def __getattr__(self, name: str) -> Any:  # delegate everything else (e.g. .optimizer, .last_epoch)
        return getattr(self.__dict__["_scheduler"], name)
