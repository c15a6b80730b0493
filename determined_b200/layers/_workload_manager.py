"""Workload manager layer: sits between the master channel and the trial controller.

Checks workload order, timestamps each workload, feeds TensorBoard, validates metrics (the searcher metric must be
present and scalar; ``None``/NaN metrics are rejected; bytes are stripped), wraps checkpoints in
``storage_mgr.store_path()`` and reports ``StorageMetadata(uuid, resources, framework, format)``, and sets
``exited_reason = USER_CANCELED`` when the trial asked to stop. Only the chief's response goes up.
Parity: reference ``harness/determined/layers/_workload_manager.py:65-324``.
"""
from __future__ import annotations

import datetime
import logging
import math
import pathlib
from typing import Any, Callable, Dict, List, Optional

import numpy as np

from .. import storage as det_storage
from .. import workload as wl
from .._env import EnvContext
from .._execution import InvalidHP

log = logging.getLogger("determined.layers")


def _now() -> str:
    return datetime.datetime.now(datetime.timezone.utc).isoformat()


def _is_scalar(v: Any) -> bool:
    return isinstance(v, (int, float, np.integer, np.floating)) and not isinstance(v, bool)


def _clean(metrics: Any) -> Any:
    if isinstance(metrics, dict):
        return {k: _clean(v) for k, v in metrics.items() if not isinstance(v, (bytes, bytearray))}
    if isinstance(metrics, (list, tuple)):
        return [_clean(v) for v in metrics]
    if isinstance(metrics, np.ndarray):
        return metrics.tolist()
    if isinstance(metrics, (np.integer,)):
        return int(metrics)
    if isinstance(metrics, (np.floating,)):
        return float(metrics)
    return metrics


class WorkloadManager(wl.Source):
    def __init__(self, env: EnvContext, workloads: wl.Stream, storage_mgr: det_storage.StorageManager,
                 tensorboard_mgr: Any = None, metric_writer: Any = None, is_chief: bool = True) -> None:
        self.env = env
        self.workloads = workloads
        self.storage_mgr = storage_mgr
        self.tensorboard_mgr = tensorboard_mgr
        self.metric_writer = metric_writer
        self.is_chief = is_chief
        self._last_step = env.initial_workload.step_id - 1 if env.initial_workload else 0

    def __iter__(self) -> wl.Stream:
        for workload, args, respond in self.workloads:
            kind = workload.kind
            if kind == wl.Workload.Kind.RUN_STEP:
                yield from self._train(workload, respond)
            elif kind == wl.Workload.Kind.COMPUTE_VALIDATION_METRICS:
                yield from self._validate(workload, respond)
            elif kind == wl.Workload.Kind.CHECKPOINT_MODEL:
                yield from self._checkpoint(workload, respond)
            elif kind == wl.Workload.Kind.TERMINATE:
                yield workload, [], lambda _r: None
                respond({})
                return
            else:
                raise AssertionError(f"unknown workload kind {kind}")

    # ---- helpers --------------------------------------------------------------------------------------------------
    def _base(self, workload: wl.Workload, start: str) -> Dict[str, Any]:
        return {"type": "WORKLOAD_COMPLETED", "workload": workload.to_json(), "start_time": start, "end_time": _now()}

    def _finish(self, respond: wl.ResponseFunc, message: Dict[str, Any], response: Dict[str, Any]) -> None:
        if response.get("stop_requested"):
            message["exited_reason"] = "USER_CANCELED"
        if response.get("invalid_hp"):
            message["exited_reason"] = "INVALID_HP"
        respond(message)

    def _train(self, workload: wl.Workload, respond: wl.ResponseFunc) -> wl.Stream:
        start = _now()
        box: Dict[str, Any] = {}
        try:
            yield workload, [], lambda r: box.setdefault("r", r)
        except InvalidHP:
            box["r"] = {"metrics": {"batch_metrics": [], "avg_metrics": {}}, "invalid_hp": True}
        r = box.get("r")
        if r is None:
            raise RuntimeError("training workload was not answered")
        if isinstance(r, wl.Skipped):
            respond(r)
            return
        metrics = _clean(r["metrics"])
        for bm in metrics.get("batch_metrics", []):
            for k, v in bm.items():
                if v is None or (_is_scalar(v) and math.isnan(float(v))):
                    raise ValueError(f"training metric {k!r} is {v}: NaN/None metrics are not allowed")
        if self.metric_writer is not None:
            self.metric_writer.on_train_step_end(workload.step_id, workload.total_batches_processed + workload.num_batches,
                                                 metrics.get("batch_metrics", []))
        if self.tensorboard_mgr is not None:
            self.tensorboard_mgr.sync()
        msg = self._base(workload, start)
        msg["metrics"] = metrics
        self._finish(respond, msg, r)

    def _validate(self, workload: wl.Workload, respond: wl.ResponseFunc) -> wl.Stream:
        start = _now()
        box: Dict[str, Any] = {}
        yield workload, [], lambda r: box.setdefault("r", r)
        r = box.get("r")
        if r is None:
            raise RuntimeError("validation workload was not answered")
        if isinstance(r, wl.Skipped):
            respond(r)
            return
        metrics = _clean(r["metrics"])
        vm = metrics.get("validation_metrics", {})
        name = self.env.experiment_config.get("searcher", {}).get("metric")
        if name:
            if name not in vm:
                raise KeyError(f"Search method is configured to use metric {name!r} but model definition returned "
                               f"validation metrics {sorted(vm)}. The metric used by the search method must be one of "
                               f"the validation metrics returned by the model definition.")
            if not _is_scalar(vm[name]):
                raise ValueError(f"searcher metric {name!r} must be a scalar, got {type(vm[name]).__name__}: {vm[name]!r}")
        for k, v in vm.items():
            if v is None or (_is_scalar(v) and math.isnan(float(v))):
                raise ValueError(f"validation metric {k!r} is {v}: NaN/None metrics are not allowed")
        if self.metric_writer is not None:
            self.metric_writer.on_validation_step_end(workload.step_id, workload.total_batches_processed, vm)
        if self.tensorboard_mgr is not None:
            self.tensorboard_mgr.sync()
        msg = self._base(workload, start)
        msg["metrics"] = metrics
        self._finish(respond, msg, r)

    def _checkpoint(self, workload: wl.Workload, respond: wl.ResponseFunc) -> wl.Stream:
        start = _now()
        box: Dict[str, Any] = {}
        if not self.is_chief:
            yield workload, [pathlib.Path("/nonexistent-non-chief")], lambda r: box.setdefault("r", r)
            respond(wl.Skipped())
            return
        with self.storage_mgr.store_path() as (storage_id, path):
            yield workload, [pathlib.Path(path)], lambda r: box.setdefault("r", r)
            r = box.get("r")
            if r is None or isinstance(r, wl.Skipped):
                raise RuntimeError("checkpoint workload was not answered by the chief")
            metadata = self.storage_mgr.store(path, r.get("framework", ""), r.get("format", ""), storage_id)
        msg = self._base(workload, start)
        msg["metrics"] = metadata.to_json()
        log.info(f"Saved trial to checkpoint {metadata.storage_id}")
        self._finish(respond, msg, {})


def build_workload_manager(env: EnvContext, workloads: wl.Stream, is_chief: bool = True) -> WorkloadManager:
    storage_cfg = env.experiment_config.get("checkpoint_storage") or {"type": "shared_fs", "host_path": "/tmp/determined-checkpoints"}
    storage_mgr = det_storage.build(storage_cfg)
    tb_mgr = writer = None
    try:
        from ..tensorboard import build as tb_build

        tb_mgr = tb_build.build_tensorboard_manager(env)
        writer = tb_build.build_metric_writer(tb_mgr) if is_chief else None
    except Exception as e:  # pragma: no cover - tensorboard is optional
        log.debug(f"tensorboard disabled: {e!r}")
    return WorkloadManager(env, workloads, storage_mgr, tb_mgr if is_chief else None, writer, is_chief)
