"""Trial start latency: keep ``torch._dynamo`` (and with it sympy, triton, FSDP — 5 to 7 s of imports) out of the start path.

``torch.optim.Optimizer.add_param_group / zero_grad / state_dict / load_state_dict`` are wrapped in
``torch._disable_dynamo``, a shim that imports ``torch._dynamo`` on first call only to mark the method as "do not trace".
Nothing here runs under ``torch.compile`` (the hot path is hand-written kernels and CUDA graphs), so the mark is
meaningless and the import is pure start-up cost — paid by every trial as soon as it constructs an optimizer. The shim
caches its result in ``fn.__dynamo_disable``; pre-seeding that attribute with the undecorated method makes it a
pass-through. A process that later does import ``torch._dynamo`` for its own reasons is unaffected apart from these four
methods being traceable.
"""
from __future__ import annotations

import sys
import time
from typing import Any, Optional

FIRST_BATCH_TS: Optional[float] = None     # wall-clock time at which this process finished its first training batch


def mark_first_batch(device: Any = None) -> None:
    """Called once by the controller after the first ``train_batch`` (synchronised, so the time is the batch's, not the
    launch's); the harness reports it in the ``startup_timing`` line."""
    global FIRST_BATCH_TS
    try:
        if device is not None and getattr(device, "type", "") == "cuda":
            import torch

            torch.cuda.synchronize(device)
    except Exception:
        pass
    FIRST_BATCH_TS = time.time()


def skip_dynamo_shims() -> int:
    """Returns how many methods were short-circuited (0 when dynamo is already loaded: nothing to save)."""
    if "torch._dynamo" in sys.modules:
        return 0
    import torch

    n = 0
    for name in ("add_param_group", "zero_grad", "state_dict", "load_state_dict"):
        fn = getattr(torch.optim.Optimizer, name, None)
        inner = getattr(fn, "__wrapped__", None)
        if inner is not None and getattr(inner, "__dynamo_disable", None) is None:
            try:
                setattr(inner, "__dynamo_disable", inner)
                n += 1
            except (AttributeError, TypeError):
                pass
    return n


def cap_cpu_threads() -> int:
    """Size this process's CPU thread pools to its share of the node: cores / slots of the agent (``DET_AGENT_SLOT_COUNT``),
    times the slots this process drives. Eight single-GPU trials each running one OpenMP team per core — the default — and
    spinning between parallel regions is what made a 0.3 s workload take 7 s on the 8-GPU box. An explicit
    ``OMP_NUM_THREADS`` wins. Returns the thread count in effect."""
    import os

    import torch

    if os.environ.get("OMP_NUM_THREADS"):
        return torch.get_num_threads()
    try:
        slots = int(os.environ.get("DET_AGENT_SLOT_COUNT", "0") or 0)
    except ValueError:
        slots = 0
    if slots <= 1:
        return torch.get_num_threads()
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    n = max(1, min(torch.get_num_threads(), cores // slots))
    torch.set_num_threads(n)
    return n
