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
