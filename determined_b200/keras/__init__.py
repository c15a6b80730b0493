"""``TFKerasTrial``: the Keras trial interface of the reference (``harness/determined/keras/_tf_keras_trial.py``,
``_tf_keras_context.py``) — ``build_model`` / ``build_training_data_loader`` / ``build_validation_data_loader`` /
``keras_callbacks``, ``context.wrap_model / wrap_optimizer / wrap_dataset / configure_fit`` — on this package's workload loop.

The controller is written against the Keras MODEL protocol, not against TensorFlow: a compiled model that offers
``train_on_batch(x, y, return_dict=True)``, ``test_on_batch(...)``, ``save_weights(path)``, ``load_weights(path)`` (every
``tf.keras`` / Keras 3 model does, on any backend). The workloads become:

* RUN_STEP — ``num_batches`` calls of ``train_on_batch`` (callbacks' ``on_train_batch_begin / end``, ``on_epoch_begin / end``
  at the epoch boundaries of ``records_per_epoch``), one metrics dictionary per batch;
* COMPUTE_VALIDATION_METRICS — ``test_on_batch`` over the validation data, batch-size-weighted means, names prefixed ``val_``
  as Keras reports them (``on_test_begin / end`` callbacks);
* CHECKPOINT_MODEL — ``save_weights`` into ``determined-keras-model-weights`` plus the loop's own position and the Python /
  NumPy RNG states; a restored trial continues at the batch it stopped at.

Data: ``(x, y)`` NumPy arrays, a ``keras.utils.Sequence``-like object (``__len__`` / ``__getitem__`` returning batches), or any
iterable of batches (a ``tf.data.Dataset`` is one). Arrays are batched to the per-slot batch size, repeated, shuffled per
epoch from the trial seed and fast-forwarded on resume.

TensorFlow and Keras are not installed in this environment, so the tests drive the controller with a NumPy model that
implements the protocol (``tests/test_keras_trial.py``). One slot per trial: gradient averaging across slots lives inside
the backend's optimizer (Horovod's ``DistributedOptimizer`` in the reference) and has no backend-neutral form — multi-slot
Keras trials are refused with a clear message. ``EstimatorTrial`` (``tf.estimator`` is TensorFlow-only and removed from
current TensorFlow) is not provided.
"""
from __future__ import annotations

import pathlib
import pickle
import random
from typing import Any, Dict, Iterator, List, Optional, Tuple

import numpy as np

from .. import _trial, _trial_controller, errors
from .._train_context import TrialContext
from ..utils import util

WEIGHTS_FILE = "determined-keras-model-weights"
STATE_FILE = "determined-keras-trial-state.pkl"


class TFKerasTrialContext(TrialContext):
    """``wrap_model`` / ``wrap_optimizer`` / ``wrap_dataset`` record what the trial built; ``configure_fit`` takes the
    ``model.fit`` arguments that still mean something here (``class_weight``, ``verbose`` are accepted and kept)."""

    def __init__(self, *args: Any, **kwargs: Any) -> None:
        super().__init__(*args, **kwargs)
        self.model: Any = None
        self.optimizers: List[Any] = []
        self.fit_args: Dict[str, Any] = {}

    def wrap_model(self, model: Any) -> Any:
        for needed in ("train_on_batch", "test_on_batch", "save_weights", "load_weights"):
            if not hasattr(model, needed):
                raise errors.InvalidModelException(f"wrap_model() wants a compiled Keras model; {type(model).__name__} has no {needed}()")
        self.model = model
        return model

    def wrap_optimizer(self, optimizer: Any) -> Any:
        self.optimizers.append(optimizer)
        return optimizer

    def wrap_dataset(self, dataset: Any) -> Any:
        return dataset

    def configure_fit(self, **kwargs: Any) -> None:
        self.fit_args.update(kwargs)


class TFKerasTrial(_trial.Trial):
    trial_context_class = TFKerasTrialContext

    def __init__(self, context: TFKerasTrialContext) -> None:
        self.context = context

    def build_model(self) -> Any:
        """Return the compiled model, passed through ``context.wrap_model``."""
        raise NotImplementedError

    def build_training_data_loader(self) -> Any:
        raise NotImplementedError

    def build_validation_data_loader(self) -> Any:
        raise NotImplementedError

    def keras_callbacks(self) -> List[Any]:
        return []

    def session_config(self) -> Any:      # TF1 session options in the reference; nothing to configure here
        return None


class _Batches:
    """Uniform view of the accepted input kinds: ``epoch(e)`` yields that epoch's batches as ``(x, y)`` or ``(x, y, w)``."""

    def __init__(self, data: Any, batch_size: int, shuffle_seed: Optional[int]) -> None:
        self.data, self.batch_size, self.seed = data, batch_size, shuffle_seed
        self.arrays = isinstance(data, (tuple, list)) and len(data) in (2, 3) and all(hasattr(a, "shape") for a in data)
        self.sequence = not self.arrays and hasattr(data, "__len__") and hasattr(data, "__getitem__")
        if not self.arrays and not self.sequence and not hasattr(data, "__iter__"):
            raise errors.InvalidDataTypeException(f"unsupported input data of type {type(data).__name__}: give (x, y) arrays, a Sequence or an iterable of batches")

    def __len__(self) -> int:                  # batches per epoch (0 = unknown: a plain iterable)
        if self.arrays:
            return len(self.data[0]) // self.batch_size if self.seed is not None else -(-len(self.data[0]) // self.batch_size)
        return len(self.data) if self.sequence else 0

    def epoch(self, e: int) -> Iterator[Tuple[Any, ...]]:
        if self.arrays:
            n = len(self.data[0])
            order = np.random.RandomState((self.seed + e) % (2 ** 32)).permutation(n) if self.seed is not None else np.arange(n)
            stop = n - n % self.batch_size if self.seed is not None else n       # training drops the ragged tail, validation keeps it
            for at in range(0, stop, self.batch_size):
                idx = order[at:at + self.batch_size]
                yield tuple(a[idx] for a in self.data)
        elif self.sequence:
            for i in range(len(self.data)):
                yield tuple(self.data[i])
            if hasattr(self.data, "on_epoch_end"):
                self.data.on_epoch_end()
        else:
            for b in self.data:
                yield tuple(b)


def _as_dict(result: Any, names: Optional[List[str]]) -> Dict[str, float]:
    if isinstance(result, dict):
        return {k: float(v) for k, v in result.items()}
    values = list(result) if isinstance(result, (list, tuple)) else [result]
    names = names or ["loss"] + [f"metric_{i}" for i in range(1, len(values))]
    return {n: float(v) for n, v in zip(names, values)}


class TFKerasTrialController(_trial_controller.CallbackTrialController):
    def __init__(self, trial_inst: TFKerasTrial, context: TFKerasTrialContext, env: Any, workloads: Any, load_path: Any = None,
                 rendezvous_info: Any = None) -> None:
        super().__init__(context, env, workloads, load_path, rendezvous_info)
        if int(env.experiment_config.get("resources", {}).get("slots_per_trial", 1)) > 1:
            raise errors.InvalidConfigurationException(
                "TFKerasTrial runs on one slot per trial here: averaging gradients across slots is done inside the backend's optimizer "
                "(Horovod's DistributedOptimizer in the reference); use PyTorchTrial for multi-slot training")
        self.trial = trial_inst
        self.model = trial_inst.build_model()
        if context.model is None:
            raise errors.InvalidModelException("build_model() must pass the model through context.wrap_model()")
        self.callbacks = list(trial_inst.keras_callbacks())
        for cb in self.callbacks:
            if hasattr(cb, "set_model"):
                cb.set_model(self.model)
        batch = context.get_per_slot_batch_size()
        self.train = _Batches(trial_inst.build_training_data_loader(), batch, env.trial_seed)
        self.validation = _Batches(trial_inst.build_validation_data_loader(), batch, None)
        if len(self.validation) == 0 and not hasattr(self.validation.data, "__iter__"):
            raise errors.InvalidDataTypeException("the validation data is empty")
        self.batches_trained = int(env.initial_workload.total_batches_processed)
        records = int(env.experiment_config.get("records_per_epoch", 0) or 0)
        self.batches_per_epoch = len(self.train) or (records // context.get_global_batch_size() if records else 0) or 0
        if load_path is not None:
            self.load(pathlib.Path(load_path))
        self._iter: Optional[Iterator[Tuple[Any, ...]]] = None
        self._epoch = -1
        self._call("on_train_begin")

    @classmethod
    def from_trial(cls, trial_inst: Any, context: Any, env: Any, workloads: Any, load_path: Any = None, rendezvous_info: Any = None,
                   **_: Any) -> "TFKerasTrialController":
        return cls(trial_inst, context, env, workloads, load_path, rendezvous_info)

    @staticmethod
    def pre_execute_hook(env: Any, dist: Any = None) -> None:
        random.seed(env.trial_seed)
        np.random.seed(env.trial_seed % (2 ** 32))
        try:        # the backend's own seed, when there is a backend
            import tensorflow as tf  # type: ignore

            tf.random.set_seed(env.trial_seed)
        except Exception:
            pass

    def _call(self, hook: str, *args: Any) -> None:
        for cb in self.callbacks:
            fn = getattr(cb, hook, None)
            if fn is not None:
                fn(*args)

    # -- the training stream: epochs back to back, fast-forwarded to the resume point -------------------------------------------
    def _next_batch(self) -> Tuple[Any, ...]:
        while True:
            if self._iter is None:
                per = self.batches_per_epoch
                self._epoch = self.batches_trained // per if per else self._epoch + 1
                self._iter = self.train.epoch(self._epoch)
                for _ in range(self.batches_trained % per if per else 0):      # resume inside an epoch
                    next(self._iter, None)
                self._call("on_epoch_begin", self._epoch, {})
            b = next(self._iter, None)
            if b is not None:
                return b
            self._call("on_epoch_end", self._epoch, {})
            self._iter = None
            if not self.batches_per_epoch and self._epoch > 10 ** 6:
                raise errors.InvalidDataTypeException("the training data yields no batches")

    def train_for_step(self, step_id: int, num_batches: int) -> Dict[str, Any]:
        names = getattr(self.model, "metrics_names", None)
        per_batch: List[Dict[str, float]] = []
        inputs = 0
        for _ in range(num_batches):
            b = self._next_batch()
            self._call("on_train_batch_begin", self.batches_trained, {})
            kwargs = {"class_weight": self.context.fit_args["class_weight"]} if self.context.fit_args.get("class_weight") else {}
            try:
                out = self.model.train_on_batch(*b, return_dict=True, **kwargs)
            except TypeError:      # older signatures without return_dict
                out = self.model.train_on_batch(*b, **kwargs)
            m = _as_dict(out, names)
            self._call("on_train_batch_end", self.batches_trained, m)
            per_batch.append(m)
            inputs += len(b[0])
            self.batches_trained += 1
            if getattr(self.model, "stop_training", False):
                self.context.set_stop_requested(True)
        return {"metrics": util.make_metrics(inputs, per_batch), "stop_requested": self.context.get_stop_requested()}

    def compute_validation_metrics(self, step_id: int) -> Dict[str, Any]:
        names = getattr(self.model, "metrics_names", None)
        self._call("on_test_begin", {})
        sums: Dict[str, float] = {}
        total = 0
        for b in self.validation.epoch(0):
            try:
                out = self.model.test_on_batch(*b, return_dict=True)
            except TypeError:
                out = self.model.test_on_batch(*b)
            n = len(b[0])
            for k, v in _as_dict(out, names).items():
                sums[k] = sums.get(k, 0.0) + v * n
            total += n
        if total == 0:
            raise errors.InvalidDataTypeException("the validation data yields no batches")
        metrics = {(k if k.startswith("val_") else "val_" + k): v / total for k, v in sums.items()}
        self._call("on_test_end", metrics)
        return {"metrics": {"validation_metrics": metrics, "num_inputs": total}, "stop_requested": self.context.get_stop_requested()}

    def save(self, path: pathlib.Path) -> None:
        path.mkdir(parents=True, exist_ok=True)
        self.model.save_weights(str(path / WEIGHTS_FILE))
        with open(path / STATE_FILE, "wb") as f:
            pickle.dump({"batches_trained": self.batches_trained, "np_rng_state": np.random.get_state(), "random_rng_state": random.getstate()}, f)

    def load(self, path: pathlib.Path) -> None:
        self.model.load_weights(str(path / WEIGHTS_FILE))
        state_file = path / STATE_FILE
        if state_file.exists():
            with open(state_file, "rb") as f:
                st = pickle.load(f)
            np.random.set_state(st["np_rng_state"])
            random.setstate(st["random_rng_state"])

    def terminate(self) -> None:
        self._call("on_train_end", {})


TFKerasTrial.trial_controller_class = TFKerasTrialController

__all__ = ["TFKerasTrial", "TFKerasTrialContext", "TFKerasTrialController"]
