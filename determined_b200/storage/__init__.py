"""Checkpoint storage managers.

Parity: reference ``common/determined_common/storage/`` — ``StorageManager`` with
``store_path()`` / ``restore_path()`` / ``delete()``, ``StorageMetadata{uuid, resources, framework, format}``,
backends ``shared_fs`` (host→container path translation), ``s3``, ``gcs``, ``hdfs``, and
``build(config)`` / ``validate_config`` (``__init__.py:31-110``, ``base.py:11-167``, ``shared.py:9-67``).
Cloud backends import their SDK lazily and raise a clear error when it is not installed.
"""
from __future__ import annotations

import contextlib
import os
import pathlib
import shutil
import tempfile
import uuid
from typing import Any, Dict, Iterator, Optional, Tuple


class StorageMetadata:
    def __init__(self, storage_id: str, resources: Dict[str, int], framework: Optional[str] = None,
                 format: Optional[str] = None) -> None:
        self.storage_id, self.resources, self.framework, self.format = storage_id, resources, framework, format

    def __json__(self) -> Dict[str, Any]:
        return {"uuid": self.storage_id, "resources": self.resources, "framework": self.framework, "format": self.format}

    to_json = __json__

    @staticmethod
    def from_json(d: Dict[str, Any]) -> "StorageMetadata":
        return StorageMetadata(d["uuid"], d.get("resources", {}), d.get("framework"), d.get("format"))

    def __repr__(self) -> str:
        return f"<storage {self.storage_id}, resources {self.resources}>"


def list_directory(root: str) -> Dict[str, int]:
    """``{relative path: size}`` for every file, ``{"dir/": 0}`` for every directory (reference ``base.py:139-167``)."""
    root_p = pathlib.Path(root)
    if not root_p.exists():
        raise FileNotFoundError(f"{root} does not exist")
    out: Dict[str, int] = {}
    for cur, dirs, files in os.walk(root):
        rel = os.path.relpath(cur, root)
        for d in dirs:
            out[(os.path.join(rel, d) if rel != "." else d) + "/"] = 0
        for f in files:
            p = os.path.join(rel, f) if rel != "." else f
            out[p] = os.path.getsize(os.path.join(cur, f))
    return out


class StorageManager:
    """Base: a local staging directory ``base_path``; remote backends upload after ``store_path`` exits and
    download before ``restore_path`` yields."""

    def __init__(self, base_path: str) -> None:
        self._base_path = base_path

    @contextlib.contextmanager
    def store_path(self, storage_id: Optional[str] = None) -> Iterator[Tuple[str, str]]:
        storage_id = storage_id or str(uuid.uuid4())
        os.makedirs(self._base_path, exist_ok=True)
        storage_dir = os.path.join(self._base_path, storage_id)
        os.makedirs(storage_dir, exist_ok=True)
        yield storage_id, storage_dir
        if not os.path.exists(storage_dir) or not os.listdir(storage_dir):
            raise RuntimeError(f"checkpoint directory {storage_dir} is empty: nothing was written")
        self.post_store_path(storage_id, storage_dir)

    def post_store_path(self, storage_id: str, storage_dir: str) -> None:
        pass

    def store(self, storage_dir: str, framework: str, fmt: str, storage_id: str) -> StorageMetadata:
        return StorageMetadata(storage_id, list_directory(storage_dir), framework, fmt)

    @contextlib.contextmanager
    def restore_path(self, metadata: StorageMetadata) -> Iterator[str]:
        raise NotImplementedError

    def delete(self, metadata: StorageMetadata) -> None:
        raise NotImplementedError


class SharedFSStorageManager(StorageManager):
    """Checkpoints on a filesystem every agent sees. ``host_path`` on the host is mounted at
    ``container_path`` inside tasks; ``storage_path`` is a sub-directory of it."""

    def __init__(self, host_path: str, container_path: Optional[str] = None, storage_path: Optional[str] = None,
                 propagation: Optional[str] = None, **_: Any) -> None:
        self.host_path = host_path
        self.container_path = container_path or host_path
        self.storage_path = storage_path
        self.propagation = propagation
        base = self.container_path
        if storage_path:
            sp = os.path.normpath(os.path.join(host_path, storage_path))
            if not sp.startswith(os.path.normpath(host_path)):
                raise ValueError("storage_path must be a subdirectory of host_path")
            base = os.path.join(self.container_path, os.path.relpath(sp, host_path))
        super().__init__(base)

    @classmethod
    def from_config(cls, config: Dict[str, Any], container_path: Optional[str] = None) -> "SharedFSStorageManager":
        allowed = {"host_path", "storage_path", "container_path", "propagation"}
        unknown = set(config) - allowed
        if unknown:
            raise TypeError(f"unknown shared_fs options: {sorted(unknown)}")
        cfg = dict(config)
        if container_path is not None:
            cfg["container_path"] = container_path
        return cls(**cfg)

    @contextlib.contextmanager
    def restore_path(self, metadata: StorageMetadata) -> Iterator[str]:
        storage_dir = os.path.join(self._base_path, metadata.storage_id)
        if not os.path.isdir(storage_dir):
            raise FileNotFoundError(f"checkpoint {metadata.storage_id} not found under {self._base_path}")
        yield storage_dir

    def delete(self, metadata: StorageMetadata) -> None:
        storage_dir = os.path.join(self._base_path, metadata.storage_id)
        if not os.path.isdir(storage_dir):
            raise FileNotFoundError(f"checkpoint {metadata.storage_id} not found under {self._base_path}")
        shutil.rmtree(storage_dir, ignore_errors=False)


class _RemoteStorageManager(StorageManager):
    """Common shape of the object-store backends: stage locally, mirror the directory to
    ``<bucket>/<uuid>/<relative path>``."""

    def __init__(self, temp_dir: Optional[str] = None) -> None:
        super().__init__(temp_dir or os.path.join(tempfile.gettempdir(), "determined-checkpoints-staging"))

    def upload(self, storage_id: str, storage_dir: str) -> None:
        raise NotImplementedError

    def download(self, metadata: StorageMetadata, storage_dir: str) -> None:
        raise NotImplementedError

    def post_store_path(self, storage_id: str, storage_dir: str) -> None:
        try:
            self.upload(storage_id, storage_dir)
        finally:
            shutil.rmtree(storage_dir, ignore_errors=True)

    @contextlib.contextmanager
    def restore_path(self, metadata: StorageMetadata) -> Iterator[str]:
        storage_dir = os.path.join(self._base_path, metadata.storage_id)
        os.makedirs(storage_dir, exist_ok=True)
        self.download(metadata, storage_dir)
        try:
            yield storage_dir
        finally:
            shutil.rmtree(storage_dir, ignore_errors=True)


class _ObjectStoreManager(_RemoteStorageManager):
    """Checkpoints as objects ``<prefix><storage_id>/<relative path>`` behind a client with put_file / get_file /
    put_bytes / delete (``utils/objstore.py``)."""

    client: Any
    prefix = ""

    def upload(self, storage_id: str, storage_dir: str) -> None:
        for rel in list_directory(storage_dir):
            key = f"{self.prefix}{storage_id}/{rel}"
            if rel.endswith("/"):
                self.client.put_bytes(key, b"")
            else:
                self.client.put_file(key, os.path.join(storage_dir, rel))

    def download(self, metadata: StorageMetadata, storage_dir: str) -> None:
        for rel in metadata.resources:
            dst = os.path.join(storage_dir, rel)
            if rel.endswith("/"):
                os.makedirs(dst, exist_ok=True)
                continue
            os.makedirs(os.path.dirname(dst) or storage_dir, exist_ok=True)
            self.client.get_file(f"{self.prefix}{metadata.storage_id}/{rel}", dst)

    def delete(self, metadata: StorageMetadata) -> None:
        for rel in metadata.resources:
            self.client.delete(f"{self.prefix}{metadata.storage_id}/{rel}")


class S3StorageManager(_ObjectStoreManager):
    """``checkpoint_storage: {type: s3, bucket, access_key?, secret_key?, endpoint_url?, prefix?}`` (reference
    ``common/determined_common/storage/s3.py``). Talks SigV4 to the endpoint itself; no boto3 needed."""

    def __init__(self, bucket: str, access_key: Optional[str] = None, secret_key: Optional[str] = None,
                 endpoint_url: Optional[str] = None, temp_dir: Optional[str] = None, prefix: Optional[str] = None,
                 region: Optional[str] = None, **_: Any) -> None:
        super().__init__(temp_dir)
        from ..utils.objstore import S3Client

        self.bucket = bucket
        self.prefix = (prefix.strip("/") + "/") if prefix else ""
        self.client = S3Client(bucket, access_key, secret_key, endpoint_url, region)


class GCSStorageManager(_ObjectStoreManager):
    """``checkpoint_storage: {type: gcs, bucket, prefix?}`` (reference ``storage/gcs.py``); credentials as in
    :class:`~determined_b200.utils.objstore.GCSClient`."""

    def __init__(self, bucket: str, temp_dir: Optional[str] = None, prefix: Optional[str] = None, **_: Any) -> None:
        super().__init__(temp_dir)
        from ..utils.objstore import GCSClient

        self.prefix = (prefix.strip("/") + "/") if prefix else ""
        self.client = GCSClient(bucket)


class HDFSStorageManager(_ObjectStoreManager):
    """``checkpoint_storage: {type: hdfs, hdfs_url, hdfs_path, user?}`` (reference ``storage/hdfs.py``) over WebHDFS."""

    def __init__(self, hdfs_url: str, hdfs_path: str, user: Optional[str] = None, temp_dir: Optional[str] = None,
                 **_: Any) -> None:
        super().__init__(temp_dir)
        from ..utils.objstore import WebHDFSClient

        self.hdfs_path = hdfs_path
        self.client = WebHDFSClient(hdfs_url, hdfs_path, user)

    def upload(self, storage_id: str, storage_dir: str) -> None:
        self.client.mkdirs(storage_id)
        for rel in list_directory(storage_dir):
            if rel.endswith("/"):
                self.client.mkdirs(f"{storage_id}/{rel}")
            else:
                self.client.put_file(f"{storage_id}/{rel}", os.path.join(storage_dir, rel))

    def delete(self, metadata: StorageMetadata) -> None:
        self.client.delete(metadata.storage_id, recursive=True)


_STORAGE_MANAGERS = {"shared_fs": SharedFSStorageManager, "s3": S3StorageManager, "gcs": GCSStorageManager,
                     "hdfs": HDFSStorageManager}


def build(config: Dict[str, Any], container_path: Optional[str] = None) -> StorageManager:
    """Build a manager from a ``checkpoint_storage`` section: retention keys (``save_*``) are not its
    business; legacy ``checkpoint_path`` / ``tensorboard_path`` fold into ``storage_path``."""
    cfg = {k: v for k, v in config.items() if not k.startswith("save_")}
    for legacy in ("checkpoint_path", "tensorboard_path"):
        if legacy in cfg:
            cfg["storage_path"] = cfg.pop(legacy)
    kind = cfg.pop("type", None)
    if kind not in _STORAGE_MANAGERS:
        raise TypeError(f"unknown checkpoint_storage type {kind!r}; expected one of {sorted(_STORAGE_MANAGERS)}")
    if kind == "shared_fs":
        return SharedFSStorageManager.from_config(cfg, container_path)
    return _STORAGE_MANAGERS[kind](**cfg)


def validate_config(config: Dict[str, Any], container_path: Optional[str] = None) -> None:
    """Prove the storage is usable before training starts: write, read back, delete a probe checkpoint."""
    mgr = build(config, container_path)
    with mgr.store_path() as (storage_id, path):
        with open(os.path.join(path, "VALIDATE.txt"), "w") as f:
            f.write(storage_id)
    meta = StorageMetadata(storage_id, {"VALIDATE.txt": len(storage_id)})
    if isinstance(mgr, SharedFSStorageManager):
        meta = StorageMetadata(storage_id, list_directory(os.path.join(mgr._base_path, storage_id)))
    with mgr.restore_path(meta) as path:
        with open(os.path.join(path, "VALIDATE.txt")) as f:
            if f.read() != storage_id:
                raise RuntimeError("checkpoint storage validation failed: read-back mismatch")
    mgr.delete(meta)
