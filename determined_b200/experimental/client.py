"""Python SDK: ``Determined`` → ``ExperimentReference`` / ``TrialReference`` / ``Checkpoint`` / ``Model``.

Parity: reference ``common/determined_common/experimental/{determined.py:11-111, experiment.py:7-110,
trial.py:7-134, checkpoint/_checkpoint.py:24-345, model.py:47-226}``, the REST helper ``api/request.py``, the
model-definition context ``context.py:71-192`` (``.detignore``, 96 MB cap) and experiment submission
``api/experiment.py:196-355``.
"""
from __future__ import annotations

import enum
import fnmatch
import json
import os
import pathlib
import shutil
import sys
import time
import urllib.error
import urllib.parse
import urllib.request
from typing import Any, Dict, Iterable, List, Optional

from .. import constants, storage
from ..utils import util


class APIError(RuntimeError):
    def __init__(self, status: int, message: str) -> None:
        super().__init__(f"{status}: {message}")
        self.status = status


class Session:
    """Tiny REST session with a bearer token cached in ``~/.config/determined_b200/auth.json``."""

    def __init__(self, master: Optional[str] = None, user: Optional[str] = None) -> None:
        master = master or util.get_default_master_address()
        if "://" not in master:
            master = "http://" + master
        self.master = master.rstrip("/")
        from ..utils import tls

        self._opener = tls.build_opener()     # https:// masters: DET_MASTER_CERT_FILE / DET_MASTER_CERT_NAME pin the certificate
        self.user = user or os.environ.get("DET_USER", "determined")
        self.token: Optional[str] = self._load_token()
        self._tried_blank_login = False

    def _store(self) -> pathlib.Path:
        return pathlib.Path(os.environ.get("DET_CONFIG_DIR", pathlib.Path.home() / ".config" / "determined_b200")) / "auth.json"

    def _load_token(self) -> Optional[str]:
        try:
            return json.loads(self._store().read_text()).get(self.master, {}).get(self.user)
        except Exception:
            return None

    def login(self, password: str = "") -> str:
        out = self.request("POST", "/api/v1/auth/login", {"username": self.user, "password": password}, auth=False)
        self.token = out["token"]
        p = self._store()
        p.parent.mkdir(parents=True, exist_ok=True)
        try:
            data = json.loads(p.read_text())
        except Exception:
            data = {}
        data.setdefault(self.master, {})[self.user] = self.token
        p.write_text(json.dumps(data))
        return self.token

    def forget_token(self) -> None:
        """Drop this (master, user) session from the token store."""
        self.token = None
        p = self._store()
        try:
            data = json.loads(p.read_text())
            data.get(self.master, {}).pop(self.user, None)
            p.write_text(json.dumps(data))
        except Exception:
            pass

    def request(self, method: str, path: str, body: Any = None, auth: bool = True, timeout: float = 60.0) -> Any:
        data = None if body is None else util.json_encode(body).encode()
        headers = {"Content-Type": "application/json"}
        if auth and self.token:
            headers["Authorization"] = f"Bearer {self.token}"
        req = urllib.request.Request(self.master + path, data=data, method=method, headers=headers)
        try:
            with self._opener.open(req, timeout=timeout) as resp:
                text = resp.read().decode()
                return json.loads(text) if text else None
        except urllib.error.HTTPError as e:
            text = e.read().decode(errors="replace")
            try:
                text = json.loads(text).get("error", text)
            except Exception:
                pass
            if e.code == 401 and auth and not self._tried_blank_login:
                # accounts start with a blank password (reference common/api/authentication.py logs the default user in the
                # same way): try that once before giving up
                self._tried_blank_login = True
                try:
                    self.login("")
                except APIError:
                    raise APIError(e.code, text) from None
                return self.request(method, path, body, auth, timeout)
            raise APIError(e.code, text) from None
        except urllib.error.URLError as e:
            raise APIError(0, f"cannot reach master at {self.master}: {e.reason}") from None

    get = lambda self, path: self.request("GET", path)  # noqa: E731
    post = lambda self, path, body=None: self.request("POST", path, body if body is not None else {})  # noqa: E731


# ---- model-definition context -------------------------------------------------------------------------------------
def read_context(directory: str, limit: int = constants.MAX_CONTEXT_SIZE) -> List[Dict[str, Any]]:
    """Walk a model-definition directory into ``[{path, content, mode}]``, honouring ``.detignore`` (gitwildmatch-ish
    patterns), skipping ``__pycache__`` / ``*.pyc``, refusing more than 96 MB."""
    root = pathlib.Path(directory)
    if not root.is_dir():
        raise ValueError(f"model definition path {directory!r} is not a directory")
    patterns: List[str] = []
    ign = root / ".detignore"
    if ign.exists():
        patterns = [l.strip() for l in ign.read_text().splitlines() if l.strip() and not l.startswith("#")]

    def ignored(rel: str, is_dir: bool) -> bool:
        name = rel.rsplit("/", 1)[-1]
        if name == "__pycache__" or name.endswith(".pyc") or name == ".detignore":
            return True
        for pat in patterns:
            p = pat.rstrip("/")
            if pat.endswith("/") and not is_dir:
                if any(fnmatch.fnmatch(part, p) for part in rel.split("/")[:-1]):
                    return True
                continue
            if fnmatch.fnmatch(rel, p) or fnmatch.fnmatch(name, p) or rel.startswith(p + "/"):
                return True
        return False

    files: List[Dict[str, Any]] = []
    total = 0
    for cur, dirs, names in os.walk(str(root)):
        rel_dir = os.path.relpath(cur, str(root))
        rel_dir = "" if rel_dir == "." else rel_dir.replace(os.sep, "/")
        dirs[:] = sorted(d for d in dirs if not ignored(f"{rel_dir}/{d}".lstrip("/"), True))
        for n in sorted(names):
            rel = f"{rel_dir}/{n}".lstrip("/")
            if ignored(rel, False):
                continue
            p = pathlib.Path(cur) / n
            if total + p.stat().st_size > limit:       # before reading it: the file alone may not fit in memory
                raise ValueError(f"model definition exceeds {util.sizeof_fmt(limit)}; use .detignore to exclude large files")
            try:
                content = p.read_text()
            except UnicodeDecodeError:
                continue  # binary artefacts do not belong in a model definition
            total += len(content)
            if total > limit:
                raise ValueError(f"model definition exceeds {util.sizeof_fmt(limit)}; use .detignore to exclude large files")
            files.append({"path": rel, "type": "file", "content": content, "mode": p.stat().st_mode & 0o777})
    return files


# ---- references --------------------------------------------------------------------------------------------------------
class CheckpointState(enum.Enum):
    UNSPECIFIED = 0
    ACTIVE = 1
    COMPLETED = 2
    ERROR = 3
    DELETED = 4


class ModelFramework(enum.Enum):
    PYTORCH = 1
    TENSORFLOW = 2


class Checkpoint:
    @staticmethod
    def parse_metadata(directory: str) -> Dict[str, Any]:
        """``metadata.json`` of a downloaded checkpoint (reference ``_checkpoint.py:parse_metadata``)."""
        with open(os.path.join(directory, "metadata.json")) as f:
            return json.load(f)

    @staticmethod
    def get_type(metadata: Dict[str, Any]) -> ModelFramework:
        """Which framework wrote the checkpoint: ``framework`` starts with ``torch`` / ``tensorflow``; without the key,
        the ``format`` decides (``pickle`` / ``cloudpickle`` are PyTorch's)."""
        fw = str(metadata.get("framework") or "")
        if fw.startswith("torch"):
            return ModelFramework.PYTORCH
        if fw.startswith("tensorflow"):
            return ModelFramework.TENSORFLOW
        if str(metadata.get("format") or "") in ("pickle", "cloudpickle"):
            return ModelFramework.PYTORCH
        if str(metadata.get("format") or "") in ("saved_model", "h5"):
            return ModelFramework.TENSORFLOW
        raise AssertionError("unknown checkpoint format: " + json.dumps({k: metadata.get(k) for k in ("framework", "format")}))

    @classmethod
    def from_json(cls, data: Dict[str, Any], session: Optional[Session] = None) -> "Checkpoint":
        """From the master's JSON of a checkpoint (``GET /api/v1/checkpoints/<uuid>`` or a model version's ``checkpoint``)."""
        return cls(session or Session(), data)

    def __init__(self, session: Session, data: Dict[str, Any]) -> None:
        self._session = session
        self.uuid = data["uuid"]
        self.experiment_id, self.trial_id = data.get("experiment_id"), data.get("trial_id")
        self.batch_number = data.get("total_batches")
        self.resources = data.get("resources") or {}
        self.framework, self.format = data.get("framework"), data.get("format")
        self.state = CheckpointState[data.get("state", "COMPLETED")]
        self.metadata = data.get("metadata") or {}
        self.hparams = data.get("hparams") or {}
        self.experiment_config = data.get("experiment_config") or {}
        self.validation = data.get("validation")
        self.determined_version = data.get("determined_version")

    def download(self, path: Optional[str] = None) -> str:
        """Copy the checkpoint out of storage into ``path`` (default ``checkpoints/<uuid>``) and write
        ``metadata.json`` next to it (the reference synthesises this file on download, too)."""
        local = pathlib.Path(path) if path else pathlib.Path("checkpoints") / self.uuid
        mgr = storage.build(self.experiment_config["checkpoint_storage"])
        meta = storage.StorageMetadata(self.uuid, self.resources, self.framework, self.format)
        if not (local / "metadata.json").exists() or not any(local.iterdir()) if local.exists() else True:
            with mgr.restore_path(meta) as src:
                if local.exists():
                    shutil.rmtree(str(local))
                shutil.copytree(src, str(local))
            self.write_metadata_file(str(local / "metadata.json"))
        return str(local)

    def write_metadata_file(self, path: str) -> None:
        with open(path, "w") as f:
            json.dump({"determined_version": self.determined_version, "framework": self.framework, "format": self.format,
                       "experiment_id": self.experiment_id, "trial_id": self.trial_id, "hparams": self.hparams,
                       "experiment_config": self.experiment_config, "metadata": self.metadata}, f, indent=2)

    def load(self, path: Optional[str] = None, **kwargs: Any) -> Any:
        """Download, then re-create the trial from ``<ckpt>/code`` and load its weights; returns the trial object
        (``trial.model`` etc.)."""
        ckpt_dir = self.download(path)
        return self.load_from_path(ckpt_dir, **kwargs)

    @staticmethod
    def load_from_path(path: str, **kwargs: Any) -> Any:
        from .. import load as det_load
        from ._native import create_trial_instance

        meta = json.loads((pathlib.Path(path) / "metadata.json").read_text())
        code = pathlib.Path(path) / "code"
        sys.path.insert(0, str(code))
        try:
            trial_class = det_load.trial_class_from_entrypoint(meta["experiment_config"]["entrypoint"])
            return create_trial_instance(trial_class, path, config=meta["experiment_config"], hparams=meta["hparams"])
        finally:
            sys.path.remove(str(code))

    def add_metadata(self, metadata: Dict[str, Any]) -> None:
        out = self._session.post(f"/api/v1/checkpoints/{self.uuid}/metadata", {"metadata": metadata})
        self.metadata = out["metadata"]

    def remove_metadata(self, keys: List[str]) -> None:
        out = self._session.post(f"/api/v1/checkpoints/{self.uuid}/metadata", {"metadata": {}, "remove": keys})
        self.metadata = out["metadata"]

    def __repr__(self) -> str:
        return f"Checkpoint(uuid={self.uuid}, trial_id={self.trial_id}, batch={self.batch_number})"


class TrialReference:
    def __init__(self, trial_id: int, session: Session) -> None:
        self.id, self._session = trial_id, session

    def describe(self) -> Dict[str, Any]:
        return self._session.get(f"/api/v1/trials/{self.id}")

    def logs(self, **filters: Any) -> List[Dict[str, Any]]:
        """The trial's log lines. Filters: ``head`` / ``tail`` / ``offset`` / ``limit``, ``agent_ids`` / ``container_ids`` /
        ``rank_ids`` / ``stdtypes`` / ``sources`` (a value or a list), ``level`` (minimum), ``timestamp_before`` /
        ``timestamp_after`` (RFC 3339), ``search``. Every line carries ``id``, its position in the log."""
        parts = []
        for k, v in filters.items():
            if v is None or v == []:
                continue
            if isinstance(v, (list, tuple, set)):
                v = ",".join(str(x) for x in v)
            parts.append(f"{k}={urllib.parse.quote(str(v), safe=',')}")
        q = "&".join(parts)
        return self._session.get(f"/api/v1/trials/{self.id}/logs" + (f"?{q}" if q else ""))

    def kill(self) -> None:
        self._session.post(f"/api/v1/trials/{self.id}/kill")

    def top_checkpoint(self, sort_by: Optional[str] = None, smaller_is_better: Optional[bool] = None) -> Checkpoint:
        return self.select_checkpoint(best=True, sort_by=sort_by, smaller_is_better=smaller_is_better)

    def select_checkpoint(self, latest: bool = False, best: bool = False, uuid: Optional[str] = None,
                          sort_by: Optional[str] = None, smaller_is_better: Optional[bool] = None) -> Checkpoint:
        if sum([int(latest), int(best), int(uuid is not None)]) != 1:
            raise AssertionError("exactly one of latest, best, or uuid must be set")
        if uuid is not None:
            return Checkpoint(self._session, self._session.get(f"/api/v1/checkpoints/{uuid}"))
        t = self.describe()
        ckpts = [c for c in t["checkpoints"] if c["state"] == "COMPLETED"]
        if not ckpts:
            raise AssertionError(f"no checkpoint found for trial {self.id}")
        if latest:
            pick = max(ckpts, key=lambda c: c["total_batches"])
        else:
            exp = self._session.get(f"/api/v1/experiments/{t['experiment_id']}")
            metric = sort_by or exp["config"]["searcher"]["metric"]
            smaller = exp["config"]["searcher"].get("smaller_is_better", True) if smaller_is_better is None else smaller_is_better
            vals = {v["total_batches"]: v["metrics"]["validation_metrics"].get(metric) for v in t.get("validations", [])}
            scored = [(vals.get(c["total_batches"]), c) for c in ckpts if vals.get(c["total_batches"]) is not None]
            if not scored:
                raise AssertionError(f"no validated checkpoint found for trial {self.id}")
            pick = (min if smaller else max)(scored, key=lambda s: s[0])[1]
        return Checkpoint(self._session, self._session.get(f"/api/v1/checkpoints/{pick['uuid']}"))


class ExperimentReference:
    def __init__(self, experiment_id: int, session: Session) -> None:
        self.id, self._session = experiment_id, session

    def describe(self, metrics: bool = False) -> Dict[str, Any]:
        return self._session.get(f"/api/v1/experiments/{self.id}" + ("?metrics=true" if metrics else ""))

    def activate(self) -> None:
        self._session.post(f"/api/v1/experiments/{self.id}/activate")

    def pause(self) -> None:
        self._session.post(f"/api/v1/experiments/{self.id}/pause")

    def cancel(self) -> None:
        self._session.post(f"/api/v1/experiments/{self.id}/cancel")

    def kill(self) -> None:
        self._session.post(f"/api/v1/experiments/{self.id}/kill")

    def archive(self) -> None:
        self._session.post(f"/api/v1/experiments/{self.id}/archive")

    def unarchive(self) -> None:
        self._session.post(f"/api/v1/experiments/{self.id}/unarchive")

    def delete(self) -> None:
        self._session.request("DELETE", f"/api/v1/experiments/{self.id}")

    def wait(self, interval: float = 0.5, timeout: Optional[float] = None) -> str:
        """Block until the experiment reaches a terminal state; returns it."""
        t0 = time.time()
        while True:
            state = self.describe()["state"]
            if state in ("COMPLETED", "CANCELED", "ERROR"):
                return state
            if timeout is not None and time.time() - t0 > timeout:
                raise TimeoutError(f"experiment {self.id} still {state} after {timeout}s")
            time.sleep(interval)

    def get_trials(self) -> List[TrialReference]:
        return [TrialReference(t["id"], self._session) for t in self.describe()["trials"]]

    def top_checkpoint(self, sort_by: Optional[str] = None, smaller_is_better: Optional[bool] = None) -> Checkpoint:
        top = self.top_n_checkpoints(1, sort_by, smaller_is_better)
        if not top:
            raise AssertionError(f"no checkpoints found for experiment {self.id}")
        return top[0]

    def top_n_checkpoints(self, limit: int, sort_by: Optional[str] = None, smaller_is_better: Optional[bool] = None) -> List[Checkpoint]:
        """The best checkpoint of each trial, best ``limit`` across the experiment."""
        exp = self.describe(metrics=True)
        metric = sort_by or exp["config"]["searcher"]["metric"]
        smaller = exp["config"]["searcher"].get("smaller_is_better", True) if smaller_is_better is None else smaller_is_better
        scored = []
        for t in exp["trials"]:
            vals = {v["total_batches"]: v["metrics"]["validation_metrics"].get(metric) for v in t.get("validations", [])}
            best = None
            for c in t["checkpoints"]:
                m = vals.get(c["total_batches"])
                if c["state"] != "COMPLETED" or m is None:
                    continue
                if best is None or (m < best[0] if smaller else m > best[0]):
                    best = (m, c)
            if best:
                scored.append(best)
        scored.sort(key=lambda s: s[0], reverse=not smaller)
        return [Checkpoint(self._session, self._session.get(f"/api/v1/checkpoints/{c['uuid']}")) for _, c in scored[:limit]]


class ModelSortBy(enum.Enum):
    """Field to sort a list of models on (reference ``experimental/model.py:10-27``)."""
    UNSPECIFIED = 0
    NAME = 1
    DESCRIPTION = 2
    CREATION_TIME = 4
    LAST_UPDATED_TIME = 5


class ModelOrderBy(enum.Enum):
    ASCENDING = 1
    ASC = 1
    DESCENDING = 2
    DESC = 2


class Model:
    def __init__(self, session: Session, data: Dict[str, Any]) -> None:
        self._session = session
        self.name, self.description = data["name"], data.get("description")
        self.metadata = data.get("metadata") or {}
        self.creation_time, self.last_updated_time = data.get("creation_time"), data.get("last_updated_time")

    def register_version(self, checkpoint_uuid: str) -> Checkpoint:
        out = self._session.post(f"/api/v1/models/{self.name}/versions", {"checkpoint_uuid": checkpoint_uuid})
        return Checkpoint(self._session, out["checkpoint"])

    def get_versions(self) -> List[Checkpoint]:
        vs = self._session.get(f"/api/v1/models/{self.name}/versions")
        return [Checkpoint(self._session, v["checkpoint"]) for v in sorted(vs, key=lambda v: -v["version"])]

    def get_version(self, version: int = 0) -> Optional[Checkpoint]:
        vs = self._session.get(f"/api/v1/models/{self.name}/versions")
        if not vs:
            return None
        pick = max(vs, key=lambda v: v["version"]) if version == 0 else next((v for v in vs if v["version"] == version), None)
        return Checkpoint(self._session, pick["checkpoint"]) if pick else None

    def add_metadata(self, metadata: Dict[str, Any]) -> None:
        self.metadata.update(metadata)
        self._session.request("PATCH", f"/api/v1/models/{self.name}", {"metadata": self.metadata})

    def remove_metadata(self, keys: List[str]) -> None:
        for k in keys:
            self.metadata.pop(k, None)
        self._session.request("PATCH", f"/api/v1/models/{self.name}", {"metadata": self.metadata})

    def set_description(self, description: str) -> None:
        self.description = description
        self._session.request("PATCH", f"/api/v1/models/{self.name}", {"description": description})

    def to_json(self) -> Dict[str, Any]:
        return {"name": self.name, "description": self.description, "metadata": self.metadata, "creation_time": self.creation_time,
                "last_updated_time": self.last_updated_time}

    @classmethod
    def from_json(cls, data: Dict[str, Any], session: Optional[Session] = None) -> "Model":
        return cls(session or Session(), data)


class Determined:
    def __init__(self, master: Optional[str] = None, user: Optional[str] = None) -> None:
        self._session = Session(master, user)

    def create_experiment(self, config: Dict[str, Any], model_dir: str, activate: bool = True, template: Optional[str] = None,
                          parent_id: Optional[int] = None, git: Optional[Dict[str, str]] = None) -> ExperimentReference:
        body: Dict[str, Any] = {"config": config, "model_definition": read_context(model_dir), "activate": activate}
        if git:
            body["git"] = git
        if template:
            body["template"] = template
        if parent_id:
            body["parent_id"] = parent_id
        out = self._session.post("/api/v1/experiments", body)
        return ExperimentReference(out["id"], self._session)

    def validate_config(self, config: Dict[str, Any]) -> Dict[str, Any]:
        return self._session.post("/api/v1/experiments", {"config": config, "validate_only": True})["config"]

    def get_experiment(self, experiment_id: int) -> ExperimentReference:
        return ExperimentReference(experiment_id, self._session)

    def get_trial(self, trial_id: int) -> TrialReference:
        return TrialReference(trial_id, self._session)

    def get_checkpoint(self, uuid: str) -> Checkpoint:
        return Checkpoint(self._session, self._session.get(f"/api/v1/checkpoints/{uuid}"))

    def create_model(self, name: str, description: str = "", metadata: Optional[Dict[str, Any]] = None) -> Model:
        return Model(self._session, self._session.post("/api/v1/models", {"name": name, "description": description, "metadata": metadata or {}}))

    def get_model(self, name: str) -> Model:
        return Model(self._session, self._session.get(f"/api/v1/models/{name}"))

    def get_models(self, sort_by: ModelSortBy = ModelSortBy.NAME, order_by: ModelOrderBy = ModelOrderBy.ASCENDING,
                   name: str = "", description: str = "") -> List[Model]:
        """Models of the registry; ``name`` / ``description`` keep the ones containing that text (case-insensitive)."""
        models = [Model(self._session, m) for m in self._session.get("/api/v1/models")]
        if name:
            models = [m for m in models if name.lower() in m.name.lower()]
        if description:
            models = [m for m in models if description.lower() in (m.description or "").lower()]
        key = {ModelSortBy.DESCRIPTION: lambda m: m.description or "", ModelSortBy.CREATION_TIME: lambda m: m.creation_time or "",
               ModelSortBy.LAST_UPDATED_TIME: lambda m: m.last_updated_time or ""}.get(sort_by, lambda m: m.name)
        return sorted(models, key=key, reverse=order_by.value == ModelOrderBy.DESCENDING.value)
