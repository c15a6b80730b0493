"""Controller hierarchy: ``TrialController`` → ``CallbackTrialController`` / ``LoopTrialController``.

Parity: reference ``harness/determined/_trial_controller.py:14-315``. A controller is the lowest harness-owned layer: it
consumes the ``workload.Stream`` produced by the layers above and runs framework code for each workload.

* :class:`CallbackTrialController` — the harness owns the loop and calls ``train_for_step`` /
  ``compute_validation_metrics`` / ``save`` / ``load``. Used by trials with re-entrant step functions (the ``no_op`` chaos
  trial of the end-to-end tests).
* :class:`LoopTrialController` — the controller owns the loop (needed for data-parallel training, where every rank walks
  the same workload sequence). Chief detection, the per-slot batch size and ``_global_barrier`` live here. The reference
  builds a ZMQ PUB/SUB + PUSH/PULL channel between chief and workers for the barrier (``:269-315``); here ranks already
  share a communicator (``determined_b200.parallel.comm``), so the barrier is one call on it.
"""
from __future__ import annotations

import abc
import logging
import pathlib
from typing import Any, Dict, Optional

from . import errors
from . import workload as wl
from ._env import EnvContext, RendezvousInfo


class TrialController(metaclass=abc.ABCMeta):
    def __init__(self, context: Any, env: EnvContext, workloads: wl.Stream, load_path: Optional[pathlib.Path] = None,
                 rendezvous_info: Optional[RendezvousInfo] = None) -> None:
        self.context = context
        self.env = env
        self.workloads = workloads
        self.load_path = load_path
        self.rendezvous_info = rendezvous_info
        self._check_if_trial_supports_configurations(env)

    @staticmethod
    @abc.abstractmethod
    def pre_execute_hook(env: EnvContext, dist: Any = None) -> Any:
        """Runs before any user code is imported or instantiated (seeding, communicator bring-up)."""

    @classmethod
    @abc.abstractmethod
    def from_trial(cls, trial_inst: Any, context: Any, env: EnvContext, workloads: wl.Stream,
                   load_path: Optional[pathlib.Path] = None, rendezvous_info: Optional[RendezvousInfo] = None,
                   **kwargs: Any) -> "TrialController":
        """Build the controller around an instantiated, framework-matched trial."""

    @classmethod
    def from_native(cls, context: Any, env: EnvContext, workloads: wl.Stream, load_path: Optional[pathlib.Path] = None,
                    rendezvous_info: Optional[RendezvousInfo] = None, **kwargs: Any) -> "TrialController":
        """Build the controller from a native-API context (a training script instead of a Trial class)."""
        raise NotImplementedError(f"{cls.__name__} does not support the native API")

    @abc.abstractmethod
    def run(self) -> None:
        """The control loop: answer every workload of the stream."""

    @staticmethod
    def supports_mixed_precision() -> bool:
        return False

    @staticmethod
    def supports_averaging_training_metrics() -> bool:
        return False

    def initialize_wrapper(self) -> None:
        pass

    def _check_if_trial_supports_configurations(self, env: EnvContext) -> None:
        cfg = env.experiment_config
        if cfg.mixed_precision_enabled() and not self.supports_mixed_precision():
            raise errors.InvalidConfigurationException(
                'mixed precision training is not supported by this trial interface; set optimizations.mixed_precision = "O0"')
        if cfg.averaging_training_metrics_enabled() and not self.supports_averaging_training_metrics():
            raise errors.InvalidConfigurationException(
                "optimizations.average_training_metrics is not supported by this trial interface")


class CallbackTrialController(TrialController):
    """Harness-owned loop over re-entrant step functions."""

    def run(self) -> None:
        for w, args, respond in self.workloads:
            try:
                if w.kind == wl.Workload.Kind.RUN_STEP:
                    response: wl.Response = self.train_for_step(w.step_id, w.num_batches)
                elif w.kind == wl.Workload.Kind.COMPUTE_VALIDATION_METRICS:
                    response = self.compute_validation_metrics(w.step_id)
                elif w.kind == wl.Workload.Kind.CHECKPOINT_MODEL:
                    if len(args) != 1 or not isinstance(args[0], pathlib.Path):
                        raise errors.InternalException("CHECKPOINT_MODEL carries exactly one path argument")
                    self.save(args[0])
                    response = {}
                elif w.kind == wl.Workload.Kind.TERMINATE:
                    self.terminate()
                    response = wl.Skipped()
                else:
                    raise errors.InternalException(f"unexpected workload {w.kind}")
            except errors.SkipWorkloadException:
                response = wl.Skipped()
            respond(response)

    @abc.abstractmethod
    def train_for_step(self, step_id: int, num_batches: int) -> Dict[str, Any]:
        """Train ``num_batches`` batches; returns ``{"metrics": make_metrics(...), "stop_requested": bool}``."""

    @abc.abstractmethod
    def compute_validation_metrics(self, step_id: int) -> Dict[str, Any]:
        """Returns ``{"metrics": {"validation_metrics": {...}, "num_inputs": n}, "stop_requested": bool}``."""

    @abc.abstractmethod
    def save(self, path: pathlib.Path) -> None:
        """Write the trial's state into the (not yet existing) directory ``path``."""

    @abc.abstractmethod
    def load(self, path: pathlib.Path) -> None:
        """Restore the state written by :meth:`save`."""

    def terminate(self) -> None:
        pass


class LoopTrialController(TrialController):
    """Controller-owned loop; every rank of a multi-slot trial runs one."""

    def __init__(self, context: Any, env: EnvContext, workloads: wl.Stream, load_path: Optional[pathlib.Path] = None,
                 rendezvous_info: Optional[RendezvousInfo] = None) -> None:
        super().__init__(context, env, workloads, load_path, rendezvous_info)
        self.comm = getattr(context, "comm", None)
        rank = self.comm.rank if self.comm is not None else 0
        self.is_chief = rank == 0
        self.batch_size = context.get_per_slot_batch_size()
        self.scheduling_unit = env.experiment_config.scheduling_unit()
        if not self.is_chief:
            logging.getLogger().setLevel(logging.DEBUG if env.experiment_config.debug_enabled() else logging.WARNING)

    def _global_barrier(self) -> None:
        """All ranks of the trial meet here (reference: ZMQ gather + broadcast between training processes)."""
        if self.comm is not None and self.comm.size > 1:
            self.comm.barrier()
