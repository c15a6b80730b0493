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


def adopt(task: dict, initial_pythonpath: str) -> None:
    env = task.get("env") or {}
    keep = set((env.get("PYTHONPATH") or "").split(":"))
    for p in initial_pythonpath.split(":"):          # paths of the trial this worker's environment was copied from
        if p and p not in keep:
            while p in sys.path:
                sys.path.remove(p)
    os.environ.update({str(k): str(v) for k, v in env.items()})
    cwd = task.get("cwd") or ""
    if cwd:
        os.chdir(cwd)
    fresh = [p for p in (env.get("PYTHONPATH") or "").split(":") if p]
    for p in reversed(fresh):
        while p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    if cwd and (not sys.path or sys.path[0] != cwd):
        sys.path.insert(0, cwd)


def main() -> int:
    initial_pp = os.environ.get("PYTHONPATH", "")
    t0 = time.time()
    try:
        cost = preload()
    except Exception as e:     # a broken preload must not eat the trial: report and fall through to a plain start
        cost = {"preload_error": repr(e)}
    ready = time.time()
    line = sys.stdin.buffer.readline()
    if not line.strip():
        return 0
    task = json.loads(line)
    adopt(task, initial_pp)
    os.environ["DET_WARM_WORKER"] = json.dumps({"preload_s": round(ready - t0, 3), "idle_s": round(time.time() - ready, 3), **cost})
    from determined_b200.exec import harness

    return harness.main()


if __name__ == "__main__":
    sys.exit(main())
