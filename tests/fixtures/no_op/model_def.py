"""A trial that trains nothing: every workload returns synthetic metrics at once, optionally after failing at random.

Same role as the reference's ``e2e_tests/tests/fixtures/no_op/model_def.py`` (a ``CallbackTrialController`` with chaos
knobs), written against this package: it lets the end-to-end tests exercise searchers, the workload sequencer, restarts
with rollback and checkpoint bookkeeping through the real master / agent / harness without any compute.

Hyperparameters: ``metrics_progression`` (decreasing | increasing | constant), ``metrics_base`` in (0, 1),
``metrics_sigma``; ``chaos_probability`` (and ``_train`` / ``_validate`` / ``_checkpoint`` overrides): chance that a workload
raises; ``fail_on_first_validation`` / ``fail_on_checkpoint_save``: message to fail with; ``request_stop``;
``training_batch_seconds`` / ``validation_seconds`` / ``save_checkpoint_seconds``: simulated durations; ``write_null``: put a NUL
byte into the trial's output (the log pipeline must carry it).
"""
import collections
import json
import pathlib
import random
import time
from typing import Any, Dict, Optional

import numpy as np

import determined as det


class NoOpTrialController(det.CallbackTrialController):
    CHECKPOINT_FILENAME = "no_op_checkpoint"

    def __init__(self, *args: Any, **kwargs: Any) -> None:
        super().__init__(*args, **kwargs)
        hp = self.env.hparams
        self.chaos = random.SystemRandom()          # not seeded: a restarted trial must not fail at the same place again
        self._batch_size = self.context.get_per_slot_batch_size()
        self.p_default = float(hp.get("chaos_probability", 0))
        self.p_train, self.p_validate, self.p_checkpoint = (hp.get("chaos_probability_train"), hp.get("chaos_probability_validate"),
                                                            hp.get("chaos_probability_checkpoint"))
        self.fail_on_first_validation = hp.get("fail_on_first_validation", "")
        self.fail_on_checkpoint_save = hp.get("fail_on_checkpoint_save", hp.get("fail_on_chechpoint_save", ""))
        self.validation_set_size = int(hp.get("validation_set_size", 32 * 32))
        self.train_batch_secs = float(hp.get("training_batch_seconds", 0))
        self.validation_secs = float(hp.get("validation_seconds", self.validation_set_size * self.train_batch_secs / self._batch_size))
        self.save_secs = float(hp.get("save_checkpoint_seconds", 0))
        self.n_train_metrics = int(hp.get("num_training_metrics", 1))
        self.n_val_metrics = int(hp.get("num_validation_metrics", 1))
        self.progression = hp.get("metrics_progression", "decreasing")
        self.base = float(hp.get("metrics_base", 0.9))
        self.sigma = float(hp.get("metrics_sigma", 0.0))
        self.request_stop = bool(hp.get("request_stop", False))
        if hp.get("write_null", False):
            print("a line with a null byte: \x00 and text after it", flush=True)
        assert self.progression in ("increasing", "decreasing", "constant") and 0 < self.base < 1 and self.sigma >= 0
        self.trained_steps: collections.Counter = collections.Counter()
        if self.load_path is not None:
            self.load(pathlib.Path(self.load_path))

    @classmethod
    def from_trial(cls, trial_inst: Any, context: Any, env: Any, workloads: Any, load_path: Any = None,
                   rendezvous_info: Any = None, **_: Any) -> "NoOpTrialController":
        return cls(context, env, workloads, load_path, rendezvous_info)

    @staticmethod
    def pre_execute_hook(env: Any, dist: Any = None) -> None:
        np.random.seed(env.trial_seed % (2 ** 32))

    def steps_trained(self) -> int:
        return sum(self.trained_steps.values())

    def current_metric(self) -> float:
        noise = float(np.random.normal(0.0, self.sigma ** 2)) if self.sigma else 0.0
        if self.progression == "constant":
            return self.base + noise
        if self.progression == "decreasing":
            return self.base ** self.steps_trained() + noise
        return 1 - self.base ** self.steps_trained() + noise

    def chaos_failure(self, probability: Optional[float]) -> None:
        p = self.p_default if probability is None else float(probability)
        if self.chaos.random() < p:
            raise Exception("CHAOS! Executing random failure.")

    def train_for_step(self, step_id: int, num_batches: int) -> Dict[str, Any]:
        if self.request_stop:
            self.context.set_stop_requested(True)
        self.chaos_failure(self.p_train)
        time.sleep(self.train_batch_secs * num_batches)
        self.trained_steps[step_id] += 1
        names = ["loss"] + [f"metric_{i}" for i in range(1, self.n_train_metrics)]
        metrics = {n: self.current_metric() for n in names}
        return {"metrics": det.util.make_metrics(self._batch_size * num_batches, [metrics] * num_batches),
                "stop_requested": self.context.get_stop_requested()}

    def compute_validation_metrics(self, step_id: int) -> Dict[str, Any]:
        if self.fail_on_first_validation:
            raise Exception(self.fail_on_first_validation)
        self.chaos_failure(self.p_validate)
        time.sleep(self.validation_secs)
        names = ["validation_error"] + [f"validation_metric_{i}" for i in range(1, self.n_val_metrics)]
        return {"metrics": {"validation_metrics": {n: self.current_metric() for n in names}, "num_inputs": self.validation_set_size},
                "stop_requested": self.context.get_stop_requested()}

    def save(self, path: pathlib.Path) -> None:
        if self.fail_on_checkpoint_save:
            raise Exception(self.fail_on_checkpoint_save)
        self.chaos_failure(self.p_checkpoint)
        time.sleep(self.save_secs)
        path.mkdir(parents=True, exist_ok=True)
        with (path / self.CHECKPOINT_FILENAME).open("w") as f:
            json.dump(self.trained_steps, f, sort_keys=True, indent=4)

    def load(self, path: pathlib.Path) -> None:
        self.chaos_failure(self.p_checkpoint)
        with (path / self.CHECKPOINT_FILENAME).open("r") as f:
            body = {int(k): int(v) for k, v in json.load(f).items()}
        assert all(k >= 0 and v >= 0 for k, v in body.items())
        self.trained_steps = collections.Counter(body)


class NoOpTrial(det.Trial):
    trial_controller_class = NoOpTrialController

    def __init__(self, context: det.TrialContext) -> None:
        self.context = context
