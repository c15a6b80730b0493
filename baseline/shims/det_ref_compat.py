"""Interpreter-level compatibility for running the UNMODIFIED reference (written for Python 3.6 / torch 1.4) on
Python 3.12 / torch 2.x: restores names that newer interpreters removed. Imported before `determined`."""
import collections
import collections.abc

for _n in ("ValuesView", "KeysView", "ItemsView", "Mapping", "MutableMapping", "Sequence", "Iterable", "Callable",
           "Iterator", "Set", "MutableSet", "MutableSequence", "Hashable", "Sized", "Container"):
    if not hasattr(collections, _n):
        setattr(collections, _n, getattr(collections.abc, _n))

import numpy as _np

for _n, _t in (("float", float), ("int", int), ("bool", bool), ("object", object), ("complex", complex), ("str", str)):
    if _n not in _np.__dict__:
        setattr(_np, _n, _t)

# torch 1.4's torch.load unpickled anything; torch >= 2.6 defaults to weights_only=True, which refuses the numpy RNG
# state every Determined checkpoint carries (reference: harness/determined/pytorch/_pytorch_trial.py:634 calls
# torch.load(path, map_location="cpu")). Restore the old default for the reference process only.
import functools as _functools

import torch as _torch

_orig_torch_load = _torch.load


@_functools.wraps(_orig_torch_load)
def _load_like_torch_1_4(*args, **kwargs):
    kwargs.setdefault("weights_only", False)
    return _orig_torch_load(*args, **kwargs)


_torch.load = _load_like_torch_1_4
