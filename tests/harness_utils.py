"""Drive a real PyTorchTrialController in-process with hand-written workload streams — no master, no sockets
(the approach of the reference's ``harness/tests/experiment/utils.py``; written anew)."""
import pathlib
from typing import Any, Callable, Dict, List, Optional, Tuple, Type

from determined_b200 import _execution, load
from determined_b200 import workload as wl


def make_controller(trial_class: Type, hparams: Dict[str, Any], workloads: wl.Stream, *, seed: int = 777,
                    load_path: Optional[pathlib.Path] = None, exp_config: Optional[Dict[str, Any]] = None,
                    initial_batches: int = 0, comm: Any = None, dist: Any = None):
    cfg = {"searcher": {"name": "single", "metric": "loss", "max_length": {"batches": 100}}}
    cfg.update(exp_config or {})
    cfg["hyperparameters"] = hparams
    env, local_dist = _execution.make_local_execution_env(True, cfg, hparams=dict(hparams), limit_gpus=1, seed=seed)
    if comm is not None:
        env.experiment_config.setdefault("resources", {})["slots_per_trial"] = comm.size
        env._per_slot_batch_size, env._global_batch_size = env._calculate_batch_sizes()
    env.initial_workload = wl.train_workload(1, 1, 1, 1, total_batches_processed=initial_batches)
    return load.load_controller_from_trial(trial_class, env, workloads, load_path, None, dist or local_dist, comm=comm)


class TrainAndValidate:
    """Workload script: ``steps`` training workloads of ``scheduling_unit`` batches, a validation after every
    ``validation_freq`` steps, optional checkpoint at the end; records what came back."""

    def __init__(self) -> None:
        self.training_metrics: List[Dict[str, Any]] = []
        self.avg_metrics: List[Dict[str, Any]] = []
        self.validation_metrics: List[Dict[str, Any]] = []
        self.checkpoint_response: Optional[Dict[str, Any]] = None

    def send(self, steps: int, validation_freq: int = 1, initial_step_id: int = 1, scheduling_unit: int = 1,
             checkpoint_dir: Optional[pathlib.Path] = None, batches_done: int = 0) -> wl.Stream:
        interceptor = wl.WorkloadResponseInterceptor()
        total = batches_done
        for step_id in range(initial_step_id, initial_step_id + steps):
            yield from interceptor.send(wl.train_workload(step_id, num_batches=scheduling_unit,
                                                          total_batches_processed=total), [])
            total += scheduling_unit
            r = interceptor.result()
            if not isinstance(r, wl.Skipped):
                m = r["metrics"]
                assert len(m["batch_metrics"]) == scheduling_unit
                self.training_metrics.extend(m["batch_metrics"])
                self.avg_metrics.append(m["avg_metrics"])
            if validation_freq and (step_id - initial_step_id + 1) % validation_freq == 0:
                yield from interceptor.send(wl.validation_workload(step_id, total_batches_processed=total), [])
                r = interceptor.result()
                if not isinstance(r, wl.Skipped):
                    self.validation_metrics.append(r["metrics"]["validation_metrics"])
        if checkpoint_dir is not None:
            yield from interceptor.send(wl.checkpoint_workload(total_batches_processed=total), [checkpoint_dir])
            r = interceptor.result()
            if not isinstance(r, wl.Skipped):
                self.checkpoint_response = r
        yield from interceptor.send(wl.terminate_workload(), [])

    def result(self) -> Tuple[List[Dict[str, Any]], List[Dict[str, Any]]]:
        return self.training_metrics, self.validation_metrics
