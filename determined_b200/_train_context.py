"""``TrialContext`` — what user code may ask about its trial, independent of the framework front-end.

Parity: reference ``harness/determined/_train_context.py:9-220`` (``get_hparam``, batch-size getters,
``get_trial_id``, ``get_experiment_config``, ``get_data_config``, ``set_stop_requested``, ``from_config``,
``distributed.get_rank/local_rank/size/num_agents``).
"""
from __future__ import annotations

from typing import Any, Dict, Optional

from ._env import DistributedContext, EnvContext


class TrialContext:
    def __init__(self, env: EnvContext, dist: Optional[DistributedContext] = None, comm: Any = None) -> None:
        self.env = env
        self.distributed = dist or DistributedContext()
        self.comm = comm          # the trial's communicator (None outside a multi-slot trial); front-ends may replace it
        self._stop_requested = False

    @classmethod
    def from_config(cls, config: Dict[str, Any], **kwargs: Any) -> "TrialContext":
        """Build a context outside of a cluster (notebooks, ``Checkpoint.load()``): hyperparameters are taken
        from constants / sampled, nothing is managed."""
        from ._execution import make_local_execution_env

        env, dist = make_local_execution_env(managed_training=False, config=config)
        return cls(env, dist, **kwargs)

    def get_experiment_config(self) -> Dict[str, Any]:
        return self.env.experiment_config

    def get_data_config(self) -> Dict[str, Any]:
        return self.env.experiment_config.get("data", {})

    def get_experiment_id(self) -> int:
        return int(self.env.det_experiment_id)

    def get_trial_id(self) -> int:
        return int(self.env.det_trial_id)

    def get_trial_seed(self) -> int:
        return self.env.trial_seed

    def get_hparams(self) -> Dict[str, Any]:
        return self.env.hparams

    def get_hparam(self, name: str) -> Any:
        if name not in self.env.hparams:
            raise KeyError(f"Could not find name '{name}' in experiment hyperparameters. Please check your "
                           f"experiment configuration 'hyperparameters' section.")
        if name == "global_batch_size":
            import logging

            logging.getLogger("determined").warning(
                "Please use get_per_slot_batch_size() or get_global_batch_size() instead of "
                "get_hparam('global_batch_size'): it is not adjusted for the number of slots.")
        return self.env.hparams[name]

    def get_per_slot_batch_size(self) -> int:
        return self.env.per_slot_batch_size

    def get_global_batch_size(self) -> int:
        return self.env.global_batch_size

    def get_initial_batch(self) -> int:
        return self.env.initial_workload.total_batches_processed

    def get_stop_requested(self) -> bool:
        return self._stop_requested

    def set_stop_requested(self, stop_requested: bool) -> None:
        """Ask the master to stop this trial gracefully after the current workload (checkpoint, then
        close). Reported as ``exited_reason = USER_CANCELED`` by the workload manager."""
        if not isinstance(stop_requested, bool):
            raise TypeError("stop_requested must be a bool")
        self._stop_requested = stop_requested


class NativeContext(TrialContext):
    """Context handed back by a framework's ``init()`` in the Native API: the user script keeps building its model
    with it and finally calls the framework's train entry, which runs ``_train_fn`` — under the harness that unwinds
    the script so the controller can take over. (Reference ``_train_context.py:158-170``.)"""

    def __init__(self, env: EnvContext, dist: Optional[DistributedContext] = None, comm: Any = None) -> None:
        super().__init__(env, dist, comm)
        self._train_fn: Any = None

    def _set_train_fn(self, train_fn: Any) -> None:
        self._train_fn = train_fn

    def _run_train_fn(self) -> None:
        if self._train_fn is None:
            raise RuntimeError("this context is not attached to a training run")
        self._train_fn()
