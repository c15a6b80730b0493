"""Warm trial worker (``python -m determined_b200.exec.warm_worker``), kept by an agent per slot.

Started with the slot's ``CUDA_VISIBLE_DEVICES`` already in place, it pays the fixed costs of a trial start ahead of time —
interpreter start, ``import torch`` and this package, CUDA context creation, loading ``libdet_kernels.so`` — and then
blocks on stdin. The agent hands it a trial as one JSON line ``{"env": {...}, "cwd": "..."}``: the worker adopts the
environment and working directory and runs the harness entry point in this same process (``csrc/core/agent.hpp``,
``ContainerManager::start``). End of input without a task means the agent is gone: exit quietly.

The reference pays container start, ``pip install`` of the harness wheel and an sshd per trial
(``master/static/srv/entrypoint.sh:1-47``, ``harness/determined/layers/_worker_process.py:113-199``).
"""
from __future__ import annotations

import json
import os
import sys
import time


def preload() -> dict:
    """Everything a trial start would otherwise wait for. Returns what it cost, in seconds."""
    cost = {}
    t = time.time()
    import numpy  # noqa: F401
    import torch

    cost["import_torch"] = round(time.time() - t, 3)
    from determined_b200.utils.fast_start import cap_cpu_threads

    cap_cpu_threads()       # also while preloading: eight workers warm up side by side
    t = time.time()
    import determined_b200  # noqa: F401
    from determined_b200 import load, pytorch, storage  # noqa: F401
    from determined_b200.exec import harness  # noqa: F401
    from determined_b200.layers import _master_client, _workload_manager  # noqa: F401

    cost["import_package"] = round(time.time() - t, 3)
    if os.environ.get("CUDA_VISIBLE_DEVICES", "") != "" and torch.cuda.is_available():
        t = time.time()
        torch.cuda.init()
        x = torch.zeros(1024, device="cuda")
        (x + 1).sum().item()
        cost["cuda_context"] = round(time.time() - t, 3)
        t = time.time()
        from determined_b200.ops import functional as F

        if F.native_available():
            cost["kernels_library"] = round(time.time() - t, 3)
    t = time.time()
    import torch._dynamo  # noqa: F401  stock optimizers' step() imports it; on the fused path it is never needed

    cost["import_dynamo"] = round(time.time() - t, 3)
    return cost


def adopt(task: dict, base_env: dict, base_path: list) -> None:
    """Become the trial: its environment (on top of this worker's own start-up environment, so nothing of the previous
    trial survives), its working directory, its import path."""
    env = task.get("env") or {}
    os.environ.clear()
    os.environ.update(base_env)
    os.environ.update({str(k): str(v) for k, v in env.items()})
    cwd = task.get("cwd") or ""
    if cwd:
        os.chdir(cwd)
    initial = set((base_env.get("PYTHONPATH") or "").split(":"))
    fresh = [p for p in (env.get("PYTHONPATH") or "").split(":") if p]
    sys.path[:] = [p for p in base_path if p not in initial or p in fresh]      # drop what only the agent's first trial needed
    for p in reversed(fresh):
        while p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    if cwd and (not sys.path or sys.path[0] != cwd):
        sys.path.insert(0, cwd)


def release(task: dict, base_modules: set) -> None:
    """After a trial that ended cleanly: forget its code and give its memory back, keep everything that made the start
    slow (interpreter, torch, this package, the CUDA context, the kernel library)."""
    import gc

    cwd = os.path.realpath((task.get("cwd") or "") or os.getcwd())
    for name, mod in list(sys.modules.items()):
        if name in base_modules:
            continue
        f = getattr(mod, "__file__", None)
        if f and os.path.realpath(f).startswith(cwd + os.sep):
            del sys.modules[name]                     # model_def and friends: the next experiment brings its own
    try:
        import faulthandler

        faulthandler.cancel_dump_traceback_later()
    except Exception:
        pass
    from determined_b200.utils import fast_start

    fast_start.FIRST_BATCH_TS = None
    gc.collect()
    try:
        import torch

        if torch.cuda.is_available() and torch.cuda.is_initialized():
            torch.cuda.synchronize()
            torch.cuda.empty_cache()
    except Exception:
        pass


def main() -> int:
    base_env = dict(os.environ)
    t0 = time.time()
    ctl = None
    try:
        fd = int(os.environ.get("DET_WARM_CONTROL_FD", "-1"))
        if fd >= 0:
            ctl = os.fdopen(fd, "w", buffering=1)
    except (OSError, ValueError):
        ctl = None
    try:
        cost = preload()
    except Exception as e:     # a broken preload must not eat the trial: report and fall through to a plain start
        cost = {"preload_error": repr(e)}
    base_path = list(sys.path)
    base_modules = set(sys.modules)
    ready = time.time()
    preload_s = round(ready - t0, 3)
    if ctl:
        ctl.write("ready\n")
    served = 0
    from determined_b200.exec import harness

    while True:
        line = sys.stdin.buffer.readline()
        if not line.strip():
            return 0                                   # the agent closed the pipe: nothing more to do
        task = json.loads(line)
        adopt(task, base_env, base_path)
        os.environ["DET_WARM_WORKER"] = json.dumps({"preload_s": preload_s, "idle_s": round(time.time() - ready, 3),
                                                    "trials_served_before": served, **cost})
        try:
            rc = harness.main()
        except SystemExit as e:
            rc = e.code if isinstance(e.code, int) else (0 if e.code is None else 1)
        except BaseException:
            import traceback

            traceback.print_exc()
            rc = 1
        sys.stdout.flush()
        sys.stderr.flush()
        served += 1
        if ctl is None or rc != 0:
            if ctl:
                ctl.write(f"done {rc}\n")
            return rc                                  # a failed trial may have left CUDA in a bad state: do not reuse
        try:
            release(task, base_modules)
        except Exception:
            ctl.write("done 0\n")
            return 0
        ready = time.time()
        ctl.write("done 0\n")


if __name__ == "__main__":
    sys.exit(main())
