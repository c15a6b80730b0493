"""Masterless ("local") execution: build the environment a trial would get from the master.

Parity: reference ``harness/determined/_execution.py:34-139`` (``_make_local_execution_env``,
``_local_execution_manager``, ``_catch_sys_exit``) and
``common/determined_common/api/experiment.py:315-355`` (client-side hyperparameter sampling used by local and
test mode).
"""
from __future__ import annotations

import contextlib
import copy
import math
import os
import pathlib
import random
import sys
from typing import Any, Dict, Iterator, Optional, Tuple


from . import constants
from . import workload as wl
from ._env import DistributedContext, EnvContext


def _merge(base: Dict[str, Any], over: Dict[str, Any]) -> Dict[str, Any]:
    out = copy.deepcopy(base)
    for k, v in (over or {}).items():
        if isinstance(v, dict) and isinstance(out.get(k), dict):
            out[k] = _merge(out[k], v)
        else:
            out[k] = copy.deepcopy(v)
    return out


def generate_random_hparam_values(hparam_def: Dict[str, Any], rng: Optional[random.Random] = None) -> Dict[str, Any]:
    """Sample one value per hyperparameter: bare values and ``const`` pass through; ``int`` is uniform on
    ``[minval, maxval)`` unless the range is a single point; ``double`` uniform; ``log`` is
    ``base ** uniform(minval, maxval)``; ``categorical`` picks from ``vals``. Nested dicts without ``type``
    are recursed into."""
    rng = rng or random.Random()
    out: Dict[str, Any] = {}
    for name, spec in (hparam_def or {}).items():
        if isinstance(spec, dict) and "type" in spec:
            t = spec["type"]
            if t == "const":
                out[name] = spec["val"]
            elif t == "int":
                lo, hi = int(spec["minval"]), int(spec["maxval"])
                out[name] = lo if hi <= lo else rng.randrange(lo, hi)
            elif t == "double":
                out[name] = rng.uniform(float(spec["minval"]), float(spec["maxval"]))
            elif t == "log":
                out[name] = math.pow(float(spec["base"]), rng.uniform(float(spec["minval"]), float(spec["maxval"])))
            elif t == "categorical":
                out[name] = rng.choice(list(spec["vals"]))
            else:
                raise ValueError(f"unknown hyperparameter type {t!r} for {name!r}")
        elif isinstance(spec, dict):
            out[name] = generate_random_hparam_values(spec, rng)
        else:
            out[name] = spec
    return out


def make_test_experiment_config(config: Dict[str, Any]) -> Dict[str, Any]:
    """Test mode (``--test``): one trial, one batch per workload, no restarts, keep no checkpoints.
    (Reference: ``api/experiment.py:258-300``.)"""
    cfg = copy.deepcopy(config)
    cfg["description"] = "[test-mode] " + str(cfg.get("description", ""))
    cfg.setdefault("checkpoint_storage", {})
    cfg["checkpoint_storage"].update({"save_experiment_best": 0, "save_trial_best": 0, "save_trial_latest": 0})
    cfg["searcher"] = {"name": "single", "metric": cfg.get("searcher", {}).get("metric", "loss"),
                       "max_length": {"batches": 1},
                       "smaller_is_better": cfg.get("searcher", {}).get("smaller_is_better", True)}
    cfg["scheduling_unit"] = 1
    cfg["max_restarts"] = 0
    cfg.pop("min_validation_period", None)
    cfg.pop("min_checkpoint_period", None)
    return cfg


def make_local_execution_env(managed_training: bool, config: Optional[Dict[str, Any]], hparams: Optional[Dict[str, Any]] = None,
                             limit_gpus: Optional[int] = None, seed: Optional[int] = None) -> Tuple[EnvContext, DistributedContext]:
    """Fake the master: defaults merged under the user's config, hyperparameters sampled, one container on
    this machine. Cluster-only sections are accepted and ignored."""
    cfg = _merge(constants.default_exp_cfg(), config or {})
    for ignored in ("bind_mounts", "environment"):
        cfg.pop(ignored, None)
    hp = hparams if hparams is not None else generate_random_hparam_values(cfg.get("hyperparameters", {}),
                                                                          random.Random(seed) if seed is not None else None)
    import torch     # not at module level: the multi-slot launcher imports this package and must start fast

    use_gpu = torch.cuda.is_available()
    n_gpus = torch.cuda.device_count() if use_gpu else 0
    if limit_gpus is not None:
        n_gpus = min(n_gpus, limit_gpus)
    dist = DistributedContext.from_environ()
    if dist.get_size() <= 1:
        cfg.setdefault("resources", {})["slots_per_trial"] = 1 if managed_training or True else 1
    else:
        cfg.setdefault("resources", {})["slots_per_trial"] = dist.get_size()
    trial_seed = seed if seed is not None else int(cfg.get("reproducibility", {}).get("experiment_seed", 0) or
                                                    random.randint(0, 1 << 31))
    env = EnvContext(master_addr="", master_port=0, container_id="local", experiment_config=cfg, hparams=hp,
                     initial_workload=wl.train_workload(1, 1, 1, cfg.get("scheduling_unit", 100)),
                     latest_checkpoint=None, use_gpu=use_gpu and n_gpus > 0,
                     container_gpus=[f"GPU-local-{i}" for i in range(n_gpus)], slot_ids=list(range(n_gpus)),
                     debug=bool(cfg.get("debug", False)), det_trial_id="1", det_experiment_id="1",
                     det_cluster_id="local", trial_seed=trial_seed, managed_training=managed_training)
    return env, dist


@contextlib.contextmanager
def local_execution_manager(new_directory: pathlib.Path) -> Iterator[None]:
    """Run with the model definition directory as CWD and first on ``sys.path`` (as in a task container)."""
    old_dir = os.getcwd()
    old_path0 = sys.path[0] if sys.path else None
    try:
        os.chdir(str(new_directory))
        sys.path.insert(0, "")
        yield
    finally:
        os.chdir(old_dir)
        try:
            sys.path.remove("")
        except ValueError:
            pass
        if old_path0 is not None and (not sys.path or sys.path[0] != old_path0):
            pass


class InvalidHP(Exception):
    """Raise from a trial's ``__init__`` / ``train_batch`` to tell the searcher this hyperparameter point is
    invalid (reported as ``exited_reason = INVALID_HP``)."""


@contextlib.contextmanager
def catch_sys_exit() -> Iterator[None]:
    try:
        yield
    except SystemExit as e:
        raise RuntimeError(f"user code called sys.exit({e.code}); trials must not exit the interpreter") from e
