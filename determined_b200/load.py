"""Locating and instantiating user trial classes.

Parity: reference ``harness/determined/load/_load_implementation.py:14-196`` (``"module:Class"``
entrypoints, reload-safe import, native ``runpy`` loading with notebook conversion) and
``load/_load_trial_controller.py:10-164`` (``pre_execute_hook`` → context → ``Trial(context)`` → controller).
"""
from __future__ import annotations

import importlib
import json
import pathlib
import runpy
import sys
import tempfile
from typing import Any, Dict, List, Optional, Tuple, Type

from . import workload as wl
from ._env import DistributedContext, EnvContext, RendezvousInfo


def trial_class_from_entrypoint(entrypoint: str) -> Type:
    """``"model_def:MyTrial"`` → the class. The module is re-imported from scratch so that a second experiment
    submitted from the same interpreter (notebook) picks up edits."""
    if ":" not in entrypoint:
        raise ValueError(f"entrypoint must look like 'module:ClassName', got {entrypoint!r}")
    module_name, class_name = entrypoint.split(":", 1)
    module_name = module_name.strip()
    if module_name.endswith(".py"):
        module_name = module_name[:-3]
    module_name = module_name.replace("/", ".")
    sys.modules.pop(module_name, None)
    importlib.invalidate_caches()
    module = importlib.import_module(module_name)
    obj: Any = module
    for part in class_name.strip().split("."):
        obj = getattr(obj, part)
    if not isinstance(obj, type):
        raise TypeError(f"entrypoint {entrypoint!r} does not name a class")
    return obj


load_trial_implementation = trial_class_from_entrypoint


def convert_notebook_to_python_script(notebook_path: pathlib.Path) -> pathlib.Path:
    """Concatenate the code cells of an ``.ipynb`` into a ``.py`` next to it (magics are dropped)."""
    nb = json.loads(pathlib.Path(notebook_path).read_text())
    lines: List[str] = []
    for cell in nb.get("cells", []):
        if cell.get("cell_type") != "code":
            continue
        src = cell.get("source", [])
        src = src if isinstance(src, list) else src.splitlines(keepends=True)
        lines += [l for l in src if not l.lstrip().startswith(("%", "!"))]
        lines.append("\n")
    out = pathlib.Path(notebook_path).with_suffix(".py")
    out.write_text("".join(lines))
    return out


class RunpyGlobals:
    """Singleton through which a natively-launched user script hands its trial class / context back."""

    _instance: Optional["RunpyGlobals"] = None

    def __init__(self, env: EnvContext, dist: DistributedContext) -> None:
        self.env, self.dist = env, dist
        self.trial_cls: Optional[Type] = None
        self.controller_cls: Optional[Type] = None
        self.context: Any = None

    def __enter__(self) -> "RunpyGlobals":
        if RunpyGlobals._instance is not None:
            raise RuntimeError("a native script is already being loaded")
        RunpyGlobals._instance = self
        return self

    def __exit__(self, *exc: Any) -> None:
        RunpyGlobals._instance = None

    @classmethod
    def is_initialized(cls) -> bool:
        return cls._instance is not None

    @classmethod
    def get_instance(cls) -> "RunpyGlobals":
        if cls._instance is None:
            raise RuntimeError("not inside a native script load")
        return cls._instance

    @classmethod
    def set_runpy_trial_result(cls, trial_cls: Type) -> None:
        cls.get_instance().trial_cls = trial_cls
        raise _StopLoadingImplementation()


    @classmethod
    def set_runpy_native_result(cls, context: Any, controller_cls: Type) -> None:
        """Native API: remember the context the script will keep using and which controller drives it; the script
        continues until it calls the framework's train entry (reference ``_load_implementation.py:117-128``)."""
        g = cls.get_instance()
        g.context, g.controller_cls = context, controller_cls


def stop_loading_implementation() -> None:
    raise _StopLoadingImplementation()


class _StopLoadingImplementation(Exception):
    pass


def load_native_implementation(env: EnvContext, dist: DistributedContext, command: List[str]) -> Any:
    """Execute the user's script (``internal.native.command``) until it either calls
    ``det.experimental.create(trial_def=...)`` (Trial API: returns the class) or, after an ``init_native(controller_cls=,
    native_context_cls=)``, reaches its train call (Native API: returns a :class:`NativeImplementation`)."""
    script = pathlib.Path(command[0])
    if script.suffix == ".ipynb":
        script = convert_notebook_to_python_script(script)
    old_argv = sys.argv
    with RunpyGlobals(env, dist) as g:
        try:
            sys.argv = [str(script)] + list(command[1:])
            runpy.run_path(str(script), run_name="__main__")
        except _StopLoadingImplementation:
            pass
        finally:
            sys.argv = old_argv
        if g.trial_cls is not None:
            return g.trial_cls
        if g.context is not None and g.controller_cls is not None:
            return NativeImplementation(g.context, g.controller_cls)
        raise RuntimeError(f"{script} finished without calling det.experimental.create(trial_def=...) or a framework init()")


class NativeImplementation:
    def __init__(self, context: Any, controller_cls: Type) -> None:
        self.context, self.controller_cls = context, controller_cls


def load_controller_from_trial(trial_class: Type, env: EnvContext, workloads: wl.Stream,
                               load_path: Optional[pathlib.Path] = None, rendezvous_info: Optional[RendezvousInfo] = None,
                               dist: Optional[DistributedContext] = None, comm: Any = None) -> Any:
    """Seed → context → user ``Trial(context)`` → controller, in the reference's order."""
    if isinstance(trial_class, NativeImplementation):
        return trial_class.controller_cls.from_native(context=trial_class.context, env=env, workloads=workloads,
                                                      load_path=load_path, rendezvous_info=rendezvous_info)
    controller_class = getattr(trial_class, "trial_controller_class", None)
    if controller_class is None:
        from .pytorch import PyTorchTrial, PyTorchTrialController

        if issubclass(trial_class, PyTorchTrial):
            controller_class = PyTorchTrialController
        else:
            raise TypeError(f"{trial_class.__name__} is not a supported Trial type")
    dist = dist or DistributedContext.from_environ()
    controller_class.pre_execute_hook(env, dist)
    context_class = getattr(trial_class, "trial_context_class")
    if comm is None:
        import torch

        from .parallel.comm import make_communicator

        dev = torch.device("cuda", dist.get_local_rank()) if (env.use_gpu and torch.cuda.is_available()) else torch.device("cpu")
        if dev.type == "cuda":
            torch.cuda.set_device(dev)
        comm = make_communicator(dev) if dist.get_size() > 1 else None
    context = context_class(env, dist, comm) if comm is not None else context_class(env, dist)
    trial_inst = trial_class(context)
    return controller_class.from_trial(trial_inst=trial_inst, context=context, env=env, workloads=workloads,
                                       load_path=load_path, rendezvous_info=rendezvous_info)


def prepare_controller_with_comm(env: EnvContext, workloads: wl.Stream, load_path: Optional[pathlib.Path],
                                 dist: DistributedContext, comm: Any) -> Any:
    """Worker-process entry: like :func:`prepare_controller` but with the trial's communicator already built (the
    workload broadcast layer needs it before the controller exists)."""
    cfg = env.experiment_config
    native = cfg.get("internal", {}) and cfg["internal"].get("native", {}).get("command")
    trial_class = load_native_implementation(env, dist, list(native)) if native else trial_class_from_entrypoint(cfg["entrypoint"])
    return load_controller_from_trial(trial_class, env, workloads, load_path, None, dist, comm=comm)


def prepare_controller(env: EnvContext, workloads: wl.Stream, load_path: Optional[pathlib.Path],
                       rendezvous_info: Optional[RendezvousInfo], dist: Optional[DistributedContext] = None) -> Any:
    """Entry used by worker processes: resolve the trial class from the experiment config and build its
    controller."""
    cfg = env.experiment_config
    native = cfg.get("internal", {}) and cfg["internal"].get("native", {}).get("command")
    dist = dist or DistributedContext.from_environ()
    if native:
        trial_class = load_native_implementation(env, dist, list(native))
    else:
        trial_class = trial_class_from_entrypoint(cfg["entrypoint"])
    return load_controller_from_trial(trial_class, env, workloads, load_path, rendezvous_info, dist)
