"""``det.Trial`` — the framework-independent base of every trial definition.

Parity: reference ``harness/determined/_trial.py``: a trial class names the controller that can run it
(``trial_controller_class``) and the context it expects (``trial_context_class``); ``from_config`` builds an instance
outside managed training (checkpoint loading, notebooks).
"""
from __future__ import annotations

import abc
from typing import Any, Dict, Optional, Type

from ._train_context import TrialContext


class Trial(metaclass=abc.ABCMeta):
    trial_controller_class: Optional[Type] = None     # a det.TrialController subclass
    trial_context_class: Type = TrialContext

    @abc.abstractmethod
    def __init__(self, context: TrialContext) -> None:
        """Framework subclasses build models / optimizers here through their context's ``wrap_*`` calls."""

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "Trial":
        """Instantiate with a context built from an experiment-config dict and no managed training loop."""
        return cls(cls.trial_context_class.from_config(config))
