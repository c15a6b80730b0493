"""Data-layer cache: build an expensive data set ONCE per (dataset_id, dataset_version), keep it on local disk, and let
every other trial / rank / restart read the cached copy — with a cluster-wide reader/writer lease from the master so
exactly one process builds while the rest wait.

Role of the reference's yogadl-backed data layer (``harness/determined/_data_layer/_data_layer.py:33-212``:
``@context.experimental.cache_train_dataset(id, version, shuffle, skip_shuffle_at_epoch_end)`` and
``cache_validation_dataset``, LFS / S3 / GCS storage, master RW coordinator ``master/internal/rw_coordinator.go``).
The reference offers it on its TensorFlow trials (tf.data); here it wraps functions that return a map-style
``torch.utils.data.Dataset``, materialises them to tensors and serves a ``TensorDataset``-like view that can also be
read as a stream positioned for this trial (``CachedDataset.stream``: resume offset, seeded per-epoch shuffle, shard of
this rank — the arguments the reference passes to yogadl at ``_data_layer.py:186-204``).

Experiment config (same keys as the reference)::

    data_layer:
      type: shared_fs | lfs
      container_storage_path: /tmp/determined-data-layer      # optional

    data_layer:
      type: s3 | gcs                       # the bucket holds the built data set, every agent keeps a local copy
      bucket: my-bucket
      bucket_directory_path: datasets/cache
      local_cache_container_path: /tmp/determined-data-layer   # optional
      access_key / secret_key / endpoint_url                    # s3 only, optional
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

    # stream position of the trial that asked for this data set; filled in by DataLayer from the trial's context
    start_offset = 0
    shuffle = False
    shuffle_seed = 0
    shard_rank = 0
    num_shards = 1
    drop_shard_remainder = False
    skip_shuffle_at_epoch_end = False

    def epoch_order(self, epoch: int) -> torch.Tensor:
        """Record ids THIS shard reads in ``epoch``: shuffle the whole set (seed + epoch, unless
        ``skip_shuffle_at_epoch_end`` pins the first permutation), then deal records round-robin to the shards."""
        n = len(self)
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(int(self.shuffle_seed) + (0 if self.skip_shuffle_at_epoch_end else epoch))
            order = torch.randperm(n, generator=g)
        else:
            order = torch.arange(n)
        if self.num_shards > 1:
            if self.drop_shard_remainder:
                order = order[: n - n % self.num_shards]
            order = order[self.shard_rank::self.num_shards]
        return order

    def stream(self, start_offset: Optional[int] = None, epochs: Optional[int] = None) -> Iterator[Any]:
        """Records from ``start_offset`` (counted in this shard's records since the start of training, so it may lie
        many epochs in) onwards; endless unless ``epochs`` bounds it."""
        offset = self.start_offset if start_offset is None else start_offset
        per_epoch = len(self.epoch_order(0))
        if per_epoch == 0:
            return
        epoch, skip = divmod(offset, per_epoch)
        last = None if epochs is None else epoch + epochs
        while last is None or epoch < last:
            for i in self.epoch_order(epoch)[skip:].tolist():
                yield self[i]
            epoch, skip = epoch + 1, 0


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
    def __init__(self, config: Optional[Dict[str, Any]] = None, master: Optional[Any] = None,
                 position: Optional[Dict[str, int]] = None) -> None:
        cfg = dict(config or {"type": "lfs"})
        kind = cfg.get("type", "lfs")
        self.remote: Any = None
        self.remote_dir = ""
        if kind in ("lfs", "shared_fs"):
            root = cfg.get("container_storage_path") or cfg.get("host_storage_path")
        elif kind in ("s3", "gcs"):
            from .utils import objstore

            if not cfg.get("bucket"):
                raise ValueError(f"data_layer.type={kind} needs `bucket`")
            self.remote = (objstore.S3Client(cfg["bucket"], cfg.get("access_key"), cfg.get("secret_key"), cfg.get("endpoint_url"))
                           if kind == "s3" else objstore.GCSClient(cfg["bucket"]))
            self.remote_dir = (cfg.get("bucket_directory_path") or "").strip("/")
            root = cfg.get("local_cache_container_path") or cfg.get("local_cache_host_path")
        else:
            raise ValueError(f"data_layer.type={kind!r}: expected one of shared_fs, lfs, s3, gcs")
        self.root = pathlib.Path(root or "/tmp/determined-data-layer")
        self.master = master
        self.position = dict(position or {})     # start_offset / shuffle_seed / shard_rank / num_shards of this trial
        self.builds = 0       # how many times THIS process built a data set (tests look at it)
        self.downloads = 0    # … and how many it fetched from the bucket

    def _path(self, dataset_id: str, version: str, split: str) -> pathlib.Path:
        safe = "".join(c if c.isalnum() or c in "-_." else "_" for c in f"{dataset_id}-{version}-{split}")
        return self.root / f"{safe}.pt"

    def _remote_key(self, path: pathlib.Path) -> str:
        return f"{self.remote_dir}/{path.name}" if self.remote_dir else path.name

    def _fetch(self, path: pathlib.Path) -> bool:
        """Local copy present (after pulling it from the bucket when there is one)?"""
        if path.exists():
            return True
        if self.remote is not None and self.remote.exists(self._remote_key(path)):
            self.root.mkdir(parents=True, exist_ok=True)
            self.remote.get_file(self._remote_key(path), str(path))
            self.downloads += 1
            return True
        return False

    def _cached(self, fn: Callable[[], Any], dataset_id: str, version: str, split: str) -> CachedDataset:
        path = self._path(dataset_id, version, split)
        lease = _Lease(self.master, path.stem, path.with_suffix(".lock"))
        local = _Lease(None, path.stem, path.with_suffix(".lock")) if self.master is not None else None
        with contextlib.ExitStack() as st:
            st.enter_context(lease.hold(write=False))
            if local is not None:       # ranks of one box share the download: the flock serialises them
                st.enter_context(local.hold(write=True))
            if self._fetch(path):
                return CachedDataset(torch.load(path))
        with lease.hold(write=True):
            if not self._fetch(path):      # somebody else may have built it while we queued for the write lease
                ds = _materialise(fn())
                self.root.mkdir(parents=True, exist_ok=True)
                tmp = path.with_suffix(f".tmp{os.getpid()}")
                torch.save(ds.columns, tmp)
                os.replace(tmp, path)
                if self.remote is not None:
                    self.remote.put_file(self._remote_key(path), str(path))
                self.builds += 1
                return ds
        return CachedDataset(torch.load(path))

    def _positioned(self, ds: CachedDataset, training: bool, shuffle: bool, skip_shuffle_at_epoch_end: bool = False) -> CachedDataset:
        pos = self.position
        ds.shuffle, ds.skip_shuffle_at_epoch_end = bool(shuffle), bool(skip_shuffle_at_epoch_end)
        ds.shuffle_seed = int(pos.get("shuffle_seed", 0))
        ds.shard_rank, ds.num_shards = int(pos.get("shard_rank", 0)), max(int(pos.get("num_shards", 1)), 1)
        ds.start_offset = int(pos.get("start_offset", 0)) if training else 0
        ds.drop_shard_remainder = training
        return ds

    def cache_train_dataset(self, dataset_id: str, dataset_version: str, shuffle: bool = False,
                            skip_shuffle_at_epoch_end: bool = False) -> Callable:
        """Decorator for a function that builds the training data set. Fed to ``det.pytorch.DataLoader`` the result is a
        plain map-style data set (the loader shuffles, shards and resumes); read through ``.stream()`` it applies
        ``shuffle`` / ``skip_shuffle_at_epoch_end`` itself, from this trial's resume offset, seed and shard."""
        def wrap(fn: Callable[..., Any]) -> Callable[..., CachedDataset]:
            @functools.wraps(fn)
            def inner(*a: Any, **k: Any) -> CachedDataset:
                return self._positioned(self._cached(lambda: fn(*a, **k), dataset_id, dataset_version, "train"), True,
                                        shuffle, skip_shuffle_at_epoch_end)
            return inner
        return wrap

    def cache_validation_dataset(self, dataset_id: str, dataset_version: str, shuffle: bool = False) -> Callable:
        def wrap(fn: Callable[..., Any]) -> Callable[..., CachedDataset]:
            @functools.wraps(fn)
            def inner(*a: Any, **k: Any) -> CachedDataset:
                return self._positioned(self._cached(lambda: fn(*a, **k), dataset_id, dataset_version, "validation"), False, shuffle)
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
    position: Dict[str, int] = {}
    if env is not None:
        try:     # reference _data_layer.py:57-69: offset = batches already trained x per-slot batch; shard = rank / size
            position = {"start_offset": int(env.initial_workload.total_batches_processed) * int(env.per_slot_batch_size),
                        "shuffle_seed": int(env.trial_seed), "shard_rank": int(context.distributed.get_rank()),
                        "num_shards": int(context.distributed.get_size())}
        except Exception:
            position = {}
    return DataLayer(cfg, master, position)
