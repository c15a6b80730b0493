"""``TensorboardManager`` tracks the tfevent files a trial writes under a local base path and mirrors new / grown
files to ``<storage>/<cluster_id>/tensorboard/experiment/<e>/trial/<t>/`` (a shared file system, or S3 / GCS objects); ``BatchMetricWriter`` writes
``Determined/<metric>`` scalars per batch and validation. TensorBoard tasks read the synced directory.
Parity: reference ``harness/determined/tensorboard/base.py:6-55``, ``build.py:10-96``,
``metric_writers/callback.py:10-55``, ``metric_writers/pytorch.py:9-42``. Shared-fs sync is a file copy; S3 / GCS sync uploads
through the clients the checkpoint storage managers use (``utils/objstore.py``)."""
from __future__ import annotations

import os
import pathlib
import shutil
from typing import Any, Dict, List, Optional

from .._env import EnvContext


def get_base_path(checkpoint_config: Dict[str, Any], manager: bool = True) -> pathlib.Path:
    base = checkpoint_config.get("base_path") or "/tmp"
    return pathlib.Path(base) / "tensorboard"


def get_sync_path(cluster_id: str, experiment_id: str, trial_id: str) -> pathlib.Path:
    return pathlib.Path(cluster_id) / "tensorboard" / "experiment" / str(experiment_id) / "trial" / str(trial_id)


class TensorboardManager:
    """``storage_root`` (shared_fs) or ``object_store`` (an S3 / GCS client plus key prefix, see :func:`object_store_for`)
    names where the event files are mirrored; with neither, ``sync`` is a no-op."""

    def __init__(self, base_path: pathlib.Path, sync_path: pathlib.Path, storage_root: Optional[pathlib.Path],
                 object_store: Optional[Any] = None, object_prefix: str = "") -> None:
        self.base_path, self.sync_path, self.storage_root = pathlib.Path(base_path), pathlib.Path(sync_path), storage_root
        self.object_store, self.object_prefix = object_store, object_prefix
        self._sizes: Dict[pathlib.Path, int] = {}
        self.base_path.mkdir(parents=True, exist_ok=True)

    def list_tfevents(self) -> List[pathlib.Path]:
        return [p for p in self.base_path.rglob("*tfevents*") if p.is_file()]

    def to_sync(self) -> List[pathlib.Path]:
        out = []
        for p in self.list_tfevents():
            size = p.stat().st_size
            if self._sizes.get(p) != size:
                self._sizes[p] = size
                out.append(p)
        return out

    def sync(self) -> None:
        if self.storage_root is None and self.object_store is None:
            return
        for p in self.to_sync():
            rel = self.sync_path / p.relative_to(self.base_path)
            if self.object_store is not None:
                self.object_store.put_file(self.object_prefix + rel.as_posix(), str(p))
                continue
            dst = self.storage_root / rel
            dst.parent.mkdir(parents=True, exist_ok=True)
            shutil.copy2(str(p), str(dst))


def object_store_for(cfg: Dict[str, Any]) -> Any:
    """(client, key prefix) for an ``s3`` / ``gcs`` storage section, else ``(None, "")`` — the same clients the checkpoint
    storage managers use (reference ``tensorboard/s3.py``, ``gcs.py``)."""
    if cfg.get("type") not in ("s3", "gcs"):
        return None, ""
    from .. import storage

    mgr = storage.build(cfg)
    return mgr.client, getattr(mgr, "prefix", "")


class BatchMetricWriter:
    """Writes per-batch training metrics and validation metrics as TensorBoard scalars."""

    def __init__(self, logdir: pathlib.Path) -> None:
        from torch.utils.tensorboard import SummaryWriter

        self.writer = SummaryWriter(log_dir=str(logdir))

    def on_train_step_end(self, step_id: int, total_batches: int, batch_metrics: List[Dict[str, Any]]) -> None:
        first = total_batches - len(batch_metrics)
        for i, bm in enumerate(batch_metrics):
            for name, value in bm.items():
                if isinstance(value, (int, float)):
                    self.writer.add_scalar(f"Determined/{name}", value, first + i)
        self.writer.flush()

    def on_validation_step_end(self, step_id: int, total_batches: int, metrics: Dict[str, Any]) -> None:
        for name, value in metrics.items():
            if isinstance(value, (int, float)):
                tag = name if name.startswith("val") else f"val_{name}"
                self.writer.add_scalar(f"Determined/{tag}", value, total_batches)
        self.writer.flush()


def build_tensorboard_manager(env: EnvContext) -> TensorboardManager:
    cfg = env.experiment_config.get("tensorboard_storage") or env.experiment_config.get("checkpoint_storage") or {}
    base = pathlib.Path(os.environ.get("DET_TENSORBOARD_BASE", "/tmp")) / f"tensorboard-{env.det_experiment_id}-{env.det_trial_id}"
    storage_root = None
    if cfg.get("type", "shared_fs") == "shared_fs" and cfg.get("host_path"):
        storage_root = pathlib.Path(cfg.get("container_path") or cfg["host_path"])
        if cfg.get("storage_path"):
            storage_root = storage_root / cfg["storage_path"]
    client, prefix = object_store_for(cfg)
    return TensorboardManager(base, get_sync_path(env.det_cluster_id, env.det_experiment_id, env.det_trial_id), storage_root, client, prefix)


def build_metric_writer(manager: TensorboardManager) -> BatchMetricWriter:
    return BatchMetricWriter(manager.base_path)
