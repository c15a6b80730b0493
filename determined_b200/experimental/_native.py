"""``det.experimental.create`` / ``test_one_batch`` / ``create_trial_instance`` — script-style entrypoints.

Parity: reference ``harness/determined/experimental/_native.py:88-399``. ``create(..., local=True,
test=True)`` runs the trial in this process for one batch: train(1) → validate → checkpoint → terminate,
through the real controller with no master, sockets or subprocesses. Cluster submission
(``local=False``) goes to the single-box master through the SDK.
"""
from __future__ import annotations

import logging
import pathlib
import tempfile
from typing import Any, Dict, List, Optional, Type

from .. import _execution, load
from .. import workload as wl
from .._env import DistributedContext, EnvContext

log = logging.getLogger("determined")


def _make_test_workloads(checkpoint_dir: pathlib.Path, config: Dict[str, Any], results: Dict[str, Any]) -> wl.Stream:
    """train one batch → validate → checkpoint → terminate, asserting each answer is well-formed."""
    interceptor = wl.WorkloadResponseInterceptor()
    log.info("Training one batch")
    yield from interceptor.send(wl.train_workload(1, num_batches=1), [])
    metrics = interceptor.metrics_result()
    batch_metrics = metrics["metrics"]["batch_metrics"]
    if len(batch_metrics) != 1 or "loss" not in batch_metrics[0]:
        raise AssertionError(f"training one batch must report exactly one batch with a 'loss': {batch_metrics}")
    results["train"] = metrics
    log.info(f"Finished training, metrics: {batch_metrics}")

    log.info("Validating one step")
    yield from interceptor.send(wl.validation_workload(1, total_batches_processed=1), [])
    validation = interceptor.metrics_result()
    v_metrics = validation["metrics"]["validation_metrics"]
    searcher_metric = config.get("searcher", {}).get("metric")
    if searcher_metric and searcher_metric not in v_metrics:
        raise AssertionError(f"searcher metric {searcher_metric!r} missing from validation metrics {sorted(v_metrics)}")
    results["validation"] = validation
    log.info(f"Finished validating, validation metrics: {v_metrics}")

    log.info(f"Saving a checkpoint to {checkpoint_dir}")
    yield from interceptor.send(wl.checkpoint_workload(1, total_batches_processed=1), [checkpoint_dir])
    results["checkpoint"] = interceptor.metrics_result()
    log.info("Finished saving a checkpoint")
    yield from interceptor.send(wl.terminate_workload(1), [])


def test_one_batch(controller_cls: Optional[Type] = None, trial_class: Optional[Type] = None,
                   config: Optional[Dict[str, Any]] = None, checkpoint_dir: Optional[str] = None) -> Dict[str, Any]:
    """Run one batch of training, one validation and one checkpoint of ``trial_class`` locally."""
    if trial_class is None:
        raise ValueError("test_one_batch needs a trial_class")
    cfg = dict(config or {})
    cfg["scheduling_unit"] = 1
    log.info("Running a minimal test experiment locally")
    results: Dict[str, Any] = {}
    with tempfile.TemporaryDirectory() as tmp:
        ckpt = pathlib.Path(checkpoint_dir) if checkpoint_dir else pathlib.Path(tmp) / "checkpoint"
        env, dist = _execution.make_local_execution_env(True, cfg, limit_gpus=1)
        workloads = _make_test_workloads(ckpt, cfg, results)
        log.info(f"Using hyperparameters: {env.hparams}")
        controller = load.load_controller_from_trial(trial_class, env, workloads, None, None, dist)
        controller.run()
        if not (ckpt / "state_dict.pth").exists():
            raise AssertionError("checkpoint workload did not write state_dict.pth")
        results["checkpoint_files"] = sorted(p.name for p in ckpt.iterdir())
    log.info("The test experiment passed")
    return results


def create_trial_instance(trial_def: Type, checkpoint_dir: str, config: Optional[Dict[str, Any]] = None,
                          hparams: Optional[Dict[str, Any]] = None) -> Any:
    """Instantiate a trial outside of training (e.g. to inspect a checkpoint): build a local environment,
    run the controller's restore path against ``checkpoint_dir`` and hand back the trial object."""
    env, dist = _execution.make_local_execution_env(False, config, hparams=hparams, limit_gpus=1)
    controller = load.load_controller_from_trial(trial_def, env, iter([]), pathlib.Path(checkpoint_dir), None, dist)
    return controller.trial


def create(trial_def: Type, config: Optional[Dict[str, Any]] = None, local: bool = False, test: bool = False,
           context_dir: str = "", command: Optional[List[str]] = None, master_url: Optional[str] = None) -> Any:
    """Create an experiment from a Trial class.

    ``local=True, test=True`` runs the one-batch smoke test in this process; ``local=False`` submits the
    experiment (config + model-definition directory) to the master; ``local=True, test=False`` is rejected
    as in the reference (``_native.py:309-313``)."""
    if load.RunpyGlobals.is_initialized():
        # we are being executed BY the harness to discover the trial class: hand it over and unwind
        load.RunpyGlobals.set_runpy_trial_result(trial_def)
    if local:
        if not test:
            raise NotImplementedError("local training is only supported in test mode: pass test=True")
        return test_one_batch(trial_class=trial_def, config=config)
    from .client import Determined

    if not context_dir:
        raise ValueError("context_dir (the model-definition directory) is required to submit to a cluster")
    cfg = dict(config or {})
    if test:
        cfg = _execution.make_test_experiment_config(cfg)
    if "entrypoint" not in cfg:
        cfg["entrypoint"] = f"{trial_def.__module__}:{trial_def.__qualname__}"
    return Determined(master_url).create_experiment(cfg, context_dir)
