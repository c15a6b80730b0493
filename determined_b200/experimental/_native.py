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


def test_one_batch(controller_cls: Optional[Type] = None, native_context_cls: Optional[Type] = None,
                   trial_class: Optional[Type] = None, config: Optional[Dict[str, Any]] = None,
                   checkpoint_dir: Optional[str] = None) -> Any:
    """Run one batch of training, one validation and one checkpoint locally. Trial API (``trial_class``): runs now and
    returns the results. Native API (``controller_cls`` + ``native_context_cls``): returns the context; the run happens
    when the script reaches its train call (reference ``_native.py:189-251``)."""
    cfg = dict(config or {})
    cfg["scheduling_unit"] = 1
    results: Dict[str, Any] = {}

    def run(build: Any) -> None:
        log.info("Running a minimal test experiment locally")
        with tempfile.TemporaryDirectory() as tmp:
            ckpt = pathlib.Path(checkpoint_dir) if checkpoint_dir else pathlib.Path(tmp) / "checkpoint"
            controller = build(_make_test_workloads(ckpt, cfg, results))
            controller.run()
            if trial_class is not None:
                # a PyTorchTrial checkpoint has state_dict.pth; other trial interfaces choose their own files, but write some
                from ..pytorch import PyTorchTrial

                if issubclass(trial_class, PyTorchTrial) and not (ckpt / "state_dict.pth").exists():
                    raise AssertionError("checkpoint workload did not write state_dict.pth")
                if not ckpt.exists() or not any(ckpt.iterdir()):
                    raise AssertionError("checkpoint workload wrote nothing")
            results["checkpoint_files"] = sorted(p.name for p in ckpt.iterdir()) if ckpt.exists() else []
        log.info("The test experiment passed")

    env, dist = _execution.make_local_execution_env(True, cfg, limit_gpus=1)
    log.info(f"Using hyperparameters: {env.hparams}")
    if native_context_cls is not None and controller_cls is not None:
        controller_cls.pre_execute_hook(env, dist)
        context = native_context_cls(env, dist)
        context._test_results = results
        context._set_train_fn(lambda: run(lambda w: controller_cls.from_native(context=context, env=env, workloads=w)))
        return context
    if trial_class is None:
        raise ValueError("test_one_batch needs a trial_class, or a controller_cls and a native_context_cls")
    run(lambda w: load.load_controller_from_trial(trial_class, env, w, None, None, dist))
    return results


def create_trial_instance(trial_def: Type, checkpoint_dir: str, config: Optional[Dict[str, Any]] = None,
                          hparams: Optional[Dict[str, Any]] = None) -> Any:
    """Instantiate a trial outside of training (e.g. to inspect a checkpoint): build a local environment,
    run the controller's restore path against ``checkpoint_dir`` and hand back the trial object."""
    env, dist = _execution.make_local_execution_env(False, config, hparams=hparams, limit_gpus=1)
    controller = load.load_controller_from_trial(trial_def, env, iter([]), pathlib.Path(checkpoint_dir), None, dist)
    return controller.trial


def init_native(trial_def: Optional[Type] = None, controller_cls: Optional[Type] = None,
                native_context_cls: Optional[Type] = None, config: Optional[Dict[str, Any]] = None, local: bool = False,
                test: bool = False, context_dir: str = "", command: Optional[List[str]] = None,
                master_url: Optional[str] = None) -> Any:
    """Entry of both user APIs (reference ``_native.py:254-292``).

    Trial API (``trial_def``): same as :func:`create`. Native API (``controller_cls`` + ``native_context_cls``): under the
    harness, build the context, register it and return it to the script — its train call then unwinds the script;
    outside the harness, submit the script (``command``) as an experiment, or with ``local=True`` run the one-batch
    smoke test."""
    if trial_def is not None:
        return create(trial_def, config=config, local=local, test=test or local, context_dir=context_dir,
                      command=command, master_url=master_url)
    if controller_cls is None or native_context_cls is None:
        raise ValueError("pass trial_def (Trial API) or controller_cls and native_context_cls (Native API)")
    if load.RunpyGlobals.is_initialized():
        g = load.RunpyGlobals.get_instance()
        controller_cls.pre_execute_hook(g.env, g.dist)
        context = native_context_cls(g.env, g.dist)
        load.RunpyGlobals.set_runpy_native_result(context, controller_cls)
        context._set_train_fn(load.stop_loading_implementation)
        return context
    if local:
        return test_one_batch(controller_cls=controller_cls, native_context_cls=native_context_cls, config=config)
    import sys

    from .client import Determined

    cfg = dict(config or {})
    if test:
        cfg = _execution.make_test_experiment_config(cfg)
    cfg.setdefault("internal", {})["native"] = {"command": list(command or sys.argv)}
    exp = Determined(master_url).create_experiment(cfg, context_dir or ".")
    return exp


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
