"""Hyperparameter-space helpers for configs written in Python (reference: ``harness/determined/_hparam.py``)."""
from typing import Any, Dict, List, Optional


def Constant(value: Any) -> Dict[str, Any]:
    return {"type": "const", "val": value}


def Integer(minval: int, maxval: int, count: Optional[int] = None) -> Dict[str, Any]:
    d: Dict[str, Any] = {"type": "int", "minval": minval, "maxval": maxval}
    if count is not None:
        d["count"] = count
    return d


def Double(minval: float, maxval: float, count: Optional[int] = None) -> Dict[str, Any]:
    d: Dict[str, Any] = {"type": "double", "minval": minval, "maxval": maxval}
    if count is not None:
        d["count"] = count
    return d


def Log(minval: float, maxval: float, base: float = 10.0, count: Optional[int] = None) -> Dict[str, Any]:
    d: Dict[str, Any] = {"type": "log", "minval": minval, "maxval": maxval, "base": base}
    if count is not None:
        d["count"] = count
    return d


def Categorical(vals: List[Any]) -> Dict[str, Any]:
    return {"type": "categorical", "vals": list(vals)}
