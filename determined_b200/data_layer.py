"""Data-layer cache: build an expensive data set ONCE per (dataset_id, dataset_version), keep it on local disk, and let
every other trial / rank / restart read the cached copy — with a cluster-wide reader/writer lease from the master so
exactly one process builds while the rest wait.

Role of the reference's yogadl-backed data layer (``harness/determined/_data_layer/_data_layer.py:33-212``:
``@context.experimental.cache_train_dataset(id, version, shuffle, skip_shuffle_at_epoch_end)`` and
``cache_validation_dataset``, LFS / S3 / GCS storage, master RW coordinator ``master/internal/rw_coordinator.go``).
The reference offers it on its TensorFlow trials (tf.data); here it wraps functions that return a map-style
``torch.utils.data.Dataset``, materialises them to tensors and serves a ``TensorDataset``-like view. Storage type
``lfs`` (local file system) is implemented; ``s3`` / ``gcs`` need SDKs that are not in this image and raise.

Experiment config (same keys as the reference)::

    data_layer:
      type: lfs
      container_storage_path: /tmp/determined-data-layer      # optional
"""
from __future__ import annotations

import contextlib
import fcntl
import functools
import os
import pathlib
import time
from typing import Any, Callable, Dict, Iterator, Optional

import torch


class CachedDataset(torch.utils.data.Dataset):
    """Tensors of a materialised data set; supports batched gathers (``det_batched``)."""

    det_batched = True

    def __init__(self, columns) -> None:
        self.columns = list(columns)

    def __len__(self) -> int:
        return int(self.columns[0].shape[0])

    def __getitem__(self, i: int):
        row = tuple(c[i] for c in self.columns)
        return row if len(row) > 1 else row[0]

    def __getitems__(self, idxs):
        idx = torch.as_tensor(idxs, dtype=torch.int64)
        row = tuple(c.index_select(0, idx) for c in self.columns)
        return row if len(row) > 1 else row[0]


def _materialise(ds: Any) -> CachedDataset:
    if isinstance(ds, CachedDataset):
        return ds
    if isinstance(ds, torch.utils.data.TensorDataset):
        return CachedDataset(ds.tensors)
    rows = [ds[i] for i in range(len(ds))]
    if not rows:
        raise ValueError("cannot cache an empty data set")
    if not isinstance(rows[0], (tuple, list)):
        rows = [(r,) for r in rows]
    cols = [torch.stack([torch.as_tensor(r[k]) for r in rows]) for k in range(len(rows[0]))]
    return CachedDataset(cols)


class _Lease:
    """Reader/writer lease: from the master when reachable, else an flock on a sibling file (single box)."""

    def __init__(self, master: Optional[Any], resource: str, lock_file: pathlib.Path) -> None:
        self.master, self.resource, self.lock_file = master, resource, lock_file

    @contextlib.contextmanager
    def hold(self, write: bool) -> Iterator[None]:
        if self.master is not None:
            verb = "write_lock" if write else "read_lock"
            lease = None
            for _ in range(240):
                out = self.master.request("POST", f"/api/v1/rw_coordinator/{self.resource}/{verb}", {"wait_ms": 20000, "ttl_ms": 600000})
                if out and out.get("granted"):
                    lease = out["lease"]
                    break
            if lease is None:
                raise TimeoutError(f"no {verb} on {self.resource} after waiting")
            try:
                yield
            finally:
                self.master.request("POST", f"/api/v1/rw_coordinator/{self.resource}/unlock", {"lease": lease})
            return
        self.lock_file.parent.mkdir(parents=True, exist_ok=True)
        with open(self.lock_file, "a+") as f:
            fcntl.flock(f, fcntl.LOCK_EX if write else fcntl.LOCK_SH)
            try:
                yield
            finally:
                fcntl.flock(f, fcntl.LOCK_UN)


class DataLayer:
    def __init__(self, config: Optional[Dict[str, Any]] = None, master: Optional[Any] = None) -> None:
        cfg = dict(config or {"type": "lfs"})
        kind = cfg.get("type", "lfs")
        if kind not in ("lfs", "shared_fs"):
            raise NotImplementedError(f"data_layer.type={kind!r} needs a cloud SDK that is not available here; use 'lfs'")
        self.root = pathlib.Path(cfg.get("container_storage_path") or cfg.get("host_storage_path") or "/tmp/determined-data-layer")
        self.master = master
        self.builds = 0    # how many times THIS process built a data set (tests look at it)

    def _path(self, dataset_id: str, version: str, split: str) -> pathlib.Path:
        safe = "".join(c if c.isalnum() or c in "-_." else "_" for c in f"{dataset_id}-{version}-{split}")
        return self.root / f"{safe}.pt"

    def _cached(self, fn: Callable[[], Any], dataset_id: str, version: str, split: str) -> CachedDataset:
        path = self._path(dataset_id, version, split)
        lease = _Lease(self.master, path.stem, path.with_suffix(".lock"))
        with lease.hold(write=False):
            if path.exists():
                return CachedDataset(torch.load(path))
        with lease.hold(write=True):
            if not path.exists():      # somebody else may have built it while we queued for the write lease
                ds = _materialise(fn())
                self.root.mkdir(parents=True, exist_ok=True)
                tmp = path.with_suffix(f".tmp{os.getpid()}")
                torch.save(ds.columns, tmp)
                os.replace(tmp, path)
                self.builds += 1
                return ds
        with lease.hold(write=False):
            return CachedDataset(torch.load(path))

    def cache_train_dataset(self, dataset_id: str, dataset_version: str, shuffle: bool = False,
                            skip_shuffle_at_epoch_end: bool = False) -> Callable:
        """Decorator for a function that builds the training data set. ``shuffle`` / ``skip_shuffle_at_epoch_end`` are
        accepted for signature parity; shuffling is the DataLoader's job here (``det.pytorch.DataLoader(shuffle=True)``
        reshuffles every epoch, and resumes mid-epoch through the skip sampler)."""
        def wrap(fn: Callable[..., Any]) -> Callable[..., CachedDataset]:
            @functools.wraps(fn)
            def inner(*a: Any, **k: Any) -> CachedDataset:
                return self._cached(lambda: fn(*a, **k), dataset_id, dataset_version, "train")
            return inner
        return wrap

    def cache_validation_dataset(self, dataset_id: str, dataset_version: str, shuffle: bool = False) -> Callable:
        def wrap(fn: Callable[..., Any]) -> Callable[..., CachedDataset]:
            @functools.wraps(fn)
            def inner(*a: Any, **k: Any) -> CachedDataset:
                return self._cached(lambda: fn(*a, **k), dataset_id, dataset_version, "validation")
            return inner
        return wrap


def from_context(context: Any) -> DataLayer:
    """Build the data layer of a trial from its experiment config and, when running under a master, its address."""
    cfg = (context.get_experiment_config() or {}).get("data_layer")
    master = None
    env = getattr(context, "env", None)
    if env is not None and getattr(env, "master_addr", None) and getattr(env, "managed_training", True) and env.master_port:
        try:
            from .layers._master_client import MasterClient

            master = MasterClient(env.master_addr, env.master_port)
            master.request("GET", "/info", retries=1)
        except Exception:
            master = None
    return DataLayer(cfg, master)
