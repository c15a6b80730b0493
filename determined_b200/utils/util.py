"""Small harness utilities: metric packaging, JSON encoding, user-code snapshotting, RNG helpers.

Parity: reference ``harness/determined/util.py:29-166`` (``make_metrics``, ``wrap_metrics``,
``json_encode``, ``write_user_code``, ``is_overridden``) and ``common/determined_common/util.py``
(``chunks``, ``sizeof_fmt``, ``preserve_random_state``).
"""
from __future__ import annotations

import contextlib
import datetime
import enum
import json
import math
import os
import pathlib
import random
import shutil
import uuid
from typing import Any, Dict, Iterator, List, Optional, Sequence, TypeVar

import numpy as np

T = TypeVar("T")


def is_overridden(full_method: Any, parent_class: Any) -> bool:
    """True if the bound method ``full_method`` is not the one defined on ``parent_class``."""
    if callable(full_method):
        return getattr(full_method, "__func__", None) is not getattr(parent_class, full_method.__name__, None)
    return full_method is not None


def _list_to_dict(list_of_dicts: List[Dict[str, Any]]) -> Dict[str, List[Any]]:
    out: Dict[str, List[Any]] = {}
    for d in list_of_dicts:
        for k, v in d.items():
            out.setdefault(k, []).append(v)
    return out


def make_metrics(num_inputs: Optional[int], batch_metrics: List[Dict[str, Any]]) -> Dict[str, Any]:
    """Package per-batch training metrics: ``avg_metrics`` is the per-name mean over batches of every metric
    that is a scalar number in every batch."""
    by_name = _list_to_dict(batch_metrics)
    avg: Dict[str, Any] = {}
    for name, values in by_name.items():
        if len(values) != len(batch_metrics):
            continue
        try:
            arr = np.asarray(values, dtype=np.float64)
        except (TypeError, ValueError):
            continue
        if arr.ndim == 1:
            avg[name] = float(np.mean(arr))
    out: Dict[str, Any] = {"batch_metrics": batch_metrics, "avg_metrics": avg}
    if num_inputs is not None:
        out["num_inputs"] = num_inputs
    return out


def wrap_metrics(metrics: Any, stop_requested: bool, invalid_hp: bool = False) -> Dict[str, Any]:
    return {"metrics": metrics, "stop_requested": stop_requested, "invalid_hp": invalid_hp}


def _json_default(o: Any) -> Any:
    if isinstance(o, (np.integer,)):
        return int(o)
    if isinstance(o, (np.floating,)):
        return _sanitize_float(float(o))
    if isinstance(o, np.bool_):
        return bool(o)
    if isinstance(o, np.ndarray):
        return o.tolist()
    if isinstance(o, (datetime.datetime, datetime.date)):
        return o.isoformat()
    if isinstance(o, enum.Enum):
        return o.name
    if isinstance(o, uuid.UUID):
        return str(o)
    if isinstance(o, (pathlib.PurePath,)):
        return str(o)
    if isinstance(o, bytes):
        return o.decode("utf-8", errors="replace")
    if hasattr(o, "__json__"):
        return o.__json__()
    if hasattr(o, "item") and callable(o.item):
        try:
            return o.item()
        except Exception:
            pass
    raise TypeError(f"{type(o).__name__} is not JSON serialisable")


def _sanitize_float(x: float) -> Any:
    return None if (math.isnan(x) or math.isinf(x)) else x


def _sanitize(o: Any) -> Any:
    if isinstance(o, float):
        return _sanitize_float(o)
    if isinstance(o, dict):
        return {str(k): _sanitize(v) for k, v in o.items()}
    if isinstance(o, (list, tuple)):
        return [_sanitize(v) for v in o]
    return o


def json_encode(obj: Any, indent: Optional[int] = None, sort_keys: bool = False) -> str:
    """JSON with numpy / datetime / enum / UUID support; NaN and ±Inf become ``null``."""
    pre = json.loads(json.dumps(obj, default=_json_default)) if not isinstance(obj, (dict, list)) else obj
    try:
        return json.dumps(_sanitize(pre), default=_json_default, indent=indent, sort_keys=sort_keys, allow_nan=False)
    except ValueError:
        # numpy floats hiding inside containers: round-trip through the default hook first
        return json.dumps(_sanitize(json.loads(json.dumps(pre, default=_json_default))), indent=indent,
                          sort_keys=sort_keys, allow_nan=False)


def write_user_code(path: pathlib.Path, on_cluster: bool = True, source: Optional[pathlib.Path] = None) -> None:
    """Snapshot the model definition (the working directory) into ``<checkpoint>/code`` so
    ``Checkpoint.load()`` can re-import the trial class later."""
    code_path = pathlib.Path(path) / "code"
    src = pathlib.Path(source) if source is not None else pathlib.Path(os.getcwd())
    if code_path.exists():
        shutil.rmtree(str(code_path))

    def ignore(_dir: str, names: List[str]) -> List[str]:
        skip = {"__pycache__", ".git", ".pytest_cache", "gpurun_out", "build", ".hypothesis", "baseline"}
        return [n for n in names if n in skip or n.endswith((".pyc", ".so", ".o"))]

    # cap the snapshot like the model-definition upload (96 MB): fall back to *.py/yaml only when larger
    total = 0
    for root, dirs, files in os.walk(str(src)):
        dirs[:] = [d for d in dirs if d not in ignore(root, dirs)]
        for f in files:
            if f in ignore(root, [f]):
                continue
            try:
                total += os.path.getsize(os.path.join(root, f))
            except OSError:
                pass
        if total > 96 * 1024 * 1024:
            break
    if total > 96 * 1024 * 1024:
        code_path.mkdir(parents=True, exist_ok=True)
        for f in src.glob("*"):
            if f.is_file() and f.suffix in (".py", ".yaml", ".yml", ".json", ".txt"):
                shutil.copy2(str(f), str(code_path / f.name))
        return
    shutil.copytree(str(src), str(code_path), ignore=ignore)


def chunks(seq: Sequence[T], n: int) -> Iterator[Sequence[T]]:
    for i in range(0, len(seq), n):
        yield seq[i:i + n]


def sizeof_fmt(val: float) -> str:
    for unit in ("", "K", "M", "G", "T", "P", "E", "Z"):
        if abs(val) < 1024.0:
            return f"{val:3.1f}{unit}B"
        val /= 1024.0
    return f"{val:.1f}YB"


@contextlib.contextmanager
def preserve_random_state() -> Iterator[None]:
    state = random.getstate()
    np_state = np.random.get_state()
    try:
        yield
    finally:
        random.setstate(state)
        np.random.set_state(np_state)


def get_default_master_address() -> str:
    return os.environ.get("DET_MASTER", os.environ.get("DET_MASTER_ADDR", "localhost:8080"))


def debug_mode() -> bool:
    return os.environ.get("DET_DEBUG", "").lower() in ("true", "1", "yes")
