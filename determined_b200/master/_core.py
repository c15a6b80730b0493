"""ctypes binding to ``libdet_core.so`` — the C++ control-plane core (NumPy-exact RNG, experiment-config
resolution/validation, searchers + simulator, workload sequencer, schedulers + fitting, LTTB).

JSON text crosses the boundary; this module turns it back into Python objects and raises on ``{"error": ...}``.
"""
from __future__ import annotations

import ctypes
import json
from pathlib import Path
from typing import Any, Dict, List, Optional

_LIB: Optional[ctypes.CDLL] = None
_LIB_PATH = Path(__file__).resolve().parent.parent / "_lib" / "libdet_core.so"


class CoreError(RuntimeError):
    pass


def load() -> ctypes.CDLL:
    global _LIB
    if _LIB is not None:
        return _LIB
    if not _LIB_PATH.exists():
        from determined_b200 import _build

        _build.build_core()
    lib = ctypes.CDLL(str(_LIB_PATH))
    vp, cp, ci, cd, ll = ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_double, ctypes.c_longlong
    sigs = {
        "det_core_free": ([vp], None),
        "det_core_nprand_new": ([ctypes.c_uint], vp), "det_core_nprand_free": ([vp], None),
        "det_core_nprand_bits32": ([vp], ctypes.c_uint), "det_core_nprand_unit_interval": ([vp], cd),
        "det_core_nprand_uniform": ([vp, cd, cd], cd), "det_core_nprand_int64n": ([vp, ll], ll),
        "det_core_nprand_int64": ([vp, ll, ll], ll), "det_core_nprand_uuid4": ([vp], vp),
        "det_core_nprand_bytes": ([vp, ctypes.c_char_p, ci], None),
        "det_core_resolve_config": ([cp], vp),
        "det_core_searcher_new": ([cp, ctypes.POINTER(vp)], vp), "det_core_searcher_free": ([vp], None),
        "det_core_searcher_initial_operations": ([vp], vp), "det_core_searcher_trial_created": ([vp, cp, ci], vp),
        "det_core_searcher_operation_completed": ([vp, ci, cp, cp], vp), "det_core_searcher_trial_closed": ([vp, cp], vp),
        "det_core_searcher_trial_exited_early": ([vp, ci, cp], vp),
        "det_core_searcher_workload_completed": ([vp, cp, cd], None), "det_core_searcher_progress": ([vp], cd),
        "det_core_searcher_events": ([vp], vp), "det_core_simulate": ([cp], vp),
        "det_core_sequencer_new": ([cp, ctypes.POINTER(vp)], vp), "det_core_sequencer_free": ([vp], None),
        "det_core_sequencer_set_trial_id": ([vp, ci], None), "det_core_sequencer_operation_requested": ([vp, cp], vp),
        "det_core_sequencer_up_to_date": ([vp], ci), "det_core_sequencer_workload": ([vp], vp),
        "det_core_sequencer_workload_completed": ([vp, cp, ci], vp),
        "det_core_sequencer_complete_cached_checkpoints": ([vp], vp), "det_core_sequencer_preclose_checkpoint": ([vp], vp),
        "det_core_sequencer_terminate_workload": ([vp], vp), "det_core_sequencer_roll_back": ([vp], ci),
        "det_core_sequencer_state": ([vp], vp),
        "det_core_schedule": ([cp], vp), "det_core_lttb": ([cp], vp), "det_core_scale_decide": ([cp], vp), "det_core_pod_spec": ([cp], vp), "det_core_cloud_command": ([cp], vp), "det_core_md5_low64": ([cp], ctypes.c_ulonglong),
    }
    for name, (args, res) in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes, fn.restype = args, res
    _LIB = lib
    return lib


def _take(ptr: int) -> Any:
    lib = load()
    if not ptr:
        raise CoreError("native call returned null")
    text = ctypes.string_at(ptr).decode()
    lib.det_core_free(ptr)
    return text


def _take_json(ptr: int) -> Any:
    out = json.loads(_take(ptr))
    if isinstance(out, dict) and set(out) == {"error"}:
        raise CoreError(out["error"])
    return out


def _enc(obj: Any) -> bytes:
    return json.dumps(obj).encode()


class NpRand:
    """NumPy ``RandomState``-compatible generator (MT19937 / RandomKit draw order)."""

    def __init__(self, seed: int) -> None:
        self._h = load().det_core_nprand_new(seed & 0xFFFFFFFF)

    def __del__(self) -> None:
        try:
            load().det_core_nprand_free(self._h)
        except Exception:
            pass

    def bits32(self) -> int:
        return load().det_core_nprand_bits32(self._h)

    def unit_interval(self) -> float:
        return load().det_core_nprand_unit_interval(self._h)

    def uniform(self, lo: float, hi: float) -> float:
        return load().det_core_nprand_uniform(self._h, lo, hi)

    def int64n(self, n: int) -> int:
        return load().det_core_nprand_int64n(self._h, n)

    def int64(self, lo: int, hi: int) -> int:
        return load().det_core_nprand_int64(self._h, lo, hi)

    def uuid4(self) -> str:
        return _take(load().det_core_nprand_uuid4(self._h))

    def bytes(self, n: int) -> bytes:
        buf = ctypes.create_string_buffer(n)
        load().det_core_nprand_bytes(self._h, buf, n)
        return buf.raw


def resolve_config(config: Dict[str, Any], master_checkpoint_storage: Optional[Dict[str, Any]] = None,
                   template: Optional[Dict[str, Any]] = None, default_seed: int = 0) -> Dict[str, Any]:
    """defaults ← master storage ← template ← user config, then validation. Raises CoreError listing every problem."""
    out = _take_json(load().det_core_resolve_config(_enc({"config": config, "master_checkpoint_storage": master_checkpoint_storage,
                                                          "template": template, "default_seed": default_seed})))
    if out["errors"]:
        raise CoreError("invalid experiment configuration: " + "; ".join(out["errors"]))
    return out["config"]


def check_config(config: Dict[str, Any]) -> List[str]:
    try:
        out = _take_json(load().det_core_resolve_config(_enc({"config": config})))
    except CoreError as e:
        return [str(e)]
    return list(out["errors"])


class Searcher:
    def __init__(self, searcher_config: Dict[str, Any], hyperparameters: Dict[str, Any], seed: int) -> None:
        err = ctypes.c_void_p()
        self._h = load().det_core_searcher_new(_enc({"searcher": searcher_config, "hyperparameters": hyperparameters, "seed": seed}),
                                               ctypes.byref(err))
        if not self._h:
            raise CoreError(_take(err.value))

    def __del__(self) -> None:
        try:
            if self._h:
                load().det_core_searcher_free(self._h)
        except Exception:
            pass

    def initial_operations(self) -> List[Dict[str, Any]]:
        return _take_json(load().det_core_searcher_initial_operations(self._h))

    def trial_created(self, create_op: Dict[str, Any], trial_id: int) -> List[Dict[str, Any]]:
        return _take_json(load().det_core_searcher_trial_created(self._h, _enc(create_op), trial_id))

    def operation_completed(self, trial_id: int, op: Dict[str, Any], metrics: Any = None) -> List[Dict[str, Any]]:
        return _take_json(load().det_core_searcher_operation_completed(self._h, trial_id, _enc(op), _enc(metrics or {})))

    def trial_closed(self, request_id: str) -> List[Dict[str, Any]]:
        return _take_json(load().det_core_searcher_trial_closed(self._h, request_id.encode()))

    def trial_exited_early(self, trial_id: int, reason: str = "ERRORED") -> List[Dict[str, Any]]:
        return _take_json(load().det_core_searcher_trial_exited_early(self._h, trial_id, reason.encode()))

    def workload_completed(self, msg: Dict[str, Any], units: float) -> None:
        load().det_core_searcher_workload_completed(self._h, _enc(msg), float(units))

    def progress(self) -> float:
        return load().det_core_searcher_progress(self._h)

    def events(self) -> List[Dict[str, Any]]:
        return _take_json(load().det_core_searcher_events(self._h))


def simulate(searcher_config: Dict[str, Any], hyperparameters: Dict[str, Any], seed: int = 0, simulation_seed: int = 0,
             validation: str = "random", random_order: bool = False) -> Dict[str, Any]:
    """Drive a searcher with synthetic validations until shutdown (``det preview-search``)."""
    return _take_json(load().det_core_simulate(_enc({"searcher": searcher_config, "hyperparameters": hyperparameters, "seed": seed,
                                                     "simulation_seed": simulation_seed, "validation": validation,
                                                     "random_order": random_order})))


class Sequencer:
    def __init__(self, config: Dict[str, Any], experiment_id: int, create_op: Dict[str, Any],
                 first_checkpoint: Optional[Dict[str, Any]] = None) -> None:
        err = ctypes.c_void_p()
        self._h = load().det_core_sequencer_new(_enc({"config": config, "experiment_id": experiment_id, "create": create_op,
                                                      "first_checkpoint": first_checkpoint}), ctypes.byref(err))
        if not self._h:
            raise CoreError(_take(err.value))

    def __del__(self) -> None:
        try:
            if self._h:
                load().det_core_sequencer_free(self._h)
        except Exception:
            pass

    def set_trial_id(self, trial_id: int) -> None:
        load().det_core_sequencer_set_trial_id(self._h, trial_id)

    def operation_requested(self, op: Dict[str, Any]) -> None:
        _take_json(load().det_core_sequencer_operation_requested(self._h, _enc(op)))

    def up_to_date(self) -> bool:
        return bool(load().det_core_sequencer_up_to_date(self._h))

    def workload(self) -> Dict[str, Any]:
        return _take_json(load().det_core_sequencer_workload(self._h))

    def workload_completed(self, msg: Dict[str, Any], is_best_validation: bool = False) -> Dict[str, Any]:
        return _take_json(load().det_core_sequencer_workload_completed(self._h, _enc(msg), int(is_best_validation)))

    def complete_cached_checkpoints(self) -> Dict[str, Any]:
        return _take_json(load().det_core_sequencer_complete_cached_checkpoints(self._h))

    def preclose_checkpoint_workload(self) -> Optional[Dict[str, Any]]:
        return _take_json(load().det_core_sequencer_preclose_checkpoint(self._h))

    def terminate_workload(self) -> Dict[str, Any]:
        return _take_json(load().det_core_sequencer_terminate_workload(self._h))

    def roll_back(self) -> int:
        return load().det_core_sequencer_roll_back(self._h)

    def state(self) -> Dict[str, Any]:
        return _take_json(load().det_core_sequencer_state(self._h))


def schedule(request: Dict[str, Any]) -> Dict[str, Any]:
    return _take_json(load().det_core_schedule(_enc(request)))


def scale_decide(config: Dict[str, Any], snapshot: Dict[str, Any]) -> Dict[str, Any]:
    """Provisioner decision for one snapshot: ``{"launch": n, "terminate": [ids], "reasons": {...}}``."""
    return _take_json(load().det_core_scale_decide(_enc({"config": config, "snapshot": snapshot})))


def pod_spec(container: Dict[str, Any], config: Dict[str, Any], cluster_id: str = "cluster", node: str = "", host_network: bool = False) -> Dict[str, Any]:
    """The pod manifest the Kubernetes back-end creates for one task container (reference: master/internal/kubernetes/spec.go)."""
    return _take_json(load().det_core_pod_spec(_enc({"container": container, "config": config, "cluster_id": cluster_id, "node": node,
                                                     "host_network": host_network})))


def cloud_command(provider: Dict[str, Any], op: str, **kw: Any) -> Dict[str, Any]:
    """What the provisioner's AWS / GCP provider would run for ``op`` (launch | list | terminate) as ``{"argv": [...]}``, or with
    ``op="parse", output=<CLI JSON>`` the instances it reads from a listing (reference: master/internal/provisioner/aws.go, gcp.go)."""
    return _take_json(load().det_core_cloud_command(_enc(dict(kw, provider=provider, op=op))))


def lttb(points: List[List[float]], threshold: int) -> List[List[float]]:
    return _take_json(load().det_core_lttb(_enc({"points": points, "threshold": threshold})))


def md5_low64(s: str) -> int:
    return load().det_core_md5_low64(s.encode())
