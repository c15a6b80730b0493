"""``determined.pytorch`` — the PyTorch Trial API (reference: ``harness/determined/pytorch/__init__.py``)."""
from ._callback import ClipGradsL2Norm, ClipGradsL2Value, PyTorchCallback
from ._context import PyTorchTrialContext
from ._data import (DataLoader, DistributedBatchSampler, RepeatBatchSampler, SkipBatchSampler, TorchData,
                    adapt_batch_sampler, data_length, to_device)
from ._lr_scheduler import LRScheduler
from ._reducer import MetricReducer, Reducer, _reduce_metrics, make_metric_reducer
from ._trial import PyTorchTrial, reset_parameters
from ._controller import PyTorchTrialController

__all__ = [
    "ClipGradsL2Norm", "ClipGradsL2Value", "PyTorchCallback", "PyTorchTrialContext", "DataLoader",
    "DistributedBatchSampler", "RepeatBatchSampler", "SkipBatchSampler", "TorchData", "adapt_batch_sampler",
    "data_length", "to_device", "LRScheduler", "MetricReducer", "Reducer", "make_metric_reducer", "PyTorchTrial",
    "PyTorchTrialController", "reset_parameters",
]
