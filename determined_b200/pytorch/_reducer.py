"""Metric reducers.

``Reducer`` (AVG/SUM/MAX/MIN) and ``_reduce_metrics`` follow reference
``harness/determined/pytorch/_reducer.py:9-43``: AVG is a mean weighted by the number of batches behind
each entry.

``MetricReducer`` is the two-method custom-reducer protocol (``accumulate`` per batch on every rank,
``cross_slot_reduce`` once on the chief over the per-rank results). In the reference it exists only for
Estimators (``estimator/_reducer.py:10-98``); the north-star asks for it on the PyTorch path, so
``PyTorchTrialContext.wrap_reducer`` / ``evaluation_reducer()`` accept it here.
"""
from __future__ import annotations

import abc
import enum
from typing import Any, Callable, List, Optional, Sequence

import numpy as np


class Reducer(enum.Enum):
    AVG = 1
    SUM = 2
    MAX = 3
    MIN = 4


def _reduce_metrics(reducer: Reducer, metrics: np.ndarray, num_batches: Optional[Sequence[int]] = None) -> float:
    if reducer == Reducer.AVG:
        if num_batches is not None and len(num_batches) != len(metrics):
            raise ValueError(f"weights ({len(num_batches)}) and metrics ({len(metrics)}) differ in length")
        return float(np.average(metrics, weights=num_batches))
    if reducer == Reducer.SUM:
        return float(np.sum(metrics))
    if reducer == Reducer.MAX:
        return float(np.max(metrics))
    if reducer == Reducer.MIN:
        return float(np.min(metrics))
    raise NotImplementedError(f"reducer {reducer} is not supported")


class MetricReducer(metaclass=abc.ABCMeta):
    """Custom distributed metric.

    * ``reset()`` — start of a validation (or training) workload;
    * ``accumulate(metric)`` — called on every rank with each batch's value; keep running state;
    * ``per_slot_reduce()`` — this rank's summary (must be picklable; tensors are moved to host);
    * ``cross_slot_reduce(per_slot_metrics)`` — on the chief, combine the list of every rank's summary into
      the final value (a scalar, or a dict of scalars when registered under several names).
    """

    def reset(self) -> None:
        pass

    @abc.abstractmethod
    def accumulate(self, metric: Any) -> Any:
        ...

    def per_slot_reduce(self) -> Any:
        return getattr(self, "_accumulated", None)

    @abc.abstractmethod
    def cross_slot_reduce(self, per_slot_metrics: List[Any]) -> Any:
        ...


class _SimpleMetricReducer(MetricReducer):
    """Adapter: a plain function over the flat list of every batch value from every rank."""

    def __init__(self, fn: Callable[[List[Any]], Any]) -> None:
        self.fn = fn
        self.updates: List[Any] = []

    def reset(self) -> None:
        self.updates = []

    def accumulate(self, metric: Any) -> Any:
        self.updates.append(metric)
        return metric

    def per_slot_reduce(self) -> Any:
        return self.updates

    def cross_slot_reduce(self, per_slot_metrics: List[Any]) -> Any:
        flat = [m for slot in per_slot_metrics for m in slot]
        return self.fn(flat)


def make_metric_reducer(fn: Callable[[List[Any]], Any]) -> MetricReducer:
    return _SimpleMetricReducer(fn)
