"""Assertion helpers that raise :class:`CheckFailedError` with a readable message.

Role of the reference's ``common/determined_common/check.py:30-321`` (used all over its harness to validate user
input and internal invariants); same function names so call sites read the same."""
from __future__ import annotations

from typing import Any, Container, Iterable, Optional, Sized, Tuple, Type, Union


class CheckFailedError(Exception):
    pass


def _fail(msg: str, reason: Optional[str]) -> None:
    raise CheckFailedError(f"CHECK FAILED! {msg}" + (f": {reason}" if reason else ""))


def true(val: Any, reason: Optional[str] = None) -> None:
    if not val:
        _fail(f"{val!r} is not truthy", reason)


def false(val: Any, reason: Optional[str] = None) -> None:
    if val:
        _fail(f"{val!r} is not falsy", reason)


def is_none(val: Any, reason: Optional[str] = None) -> None:
    if val is not None:
        _fail(f"{val!r} is not None", reason)


def is_not_none(val: Any, reason: Optional[str] = None) -> None:
    if val is None:
        _fail("value is None", reason)


def eq(a: Any, b: Any, reason: Optional[str] = None) -> None:
    if a != b:
        _fail(f"{a!r} does not equal {b!r}", reason)


def not_eq(a: Any, b: Any, reason: Optional[str] = None) -> None:
    if a == b:
        _fail(f"{a!r} equals {b!r}", reason)


def gt(a: Any, b: Any, reason: Optional[str] = None) -> None:
    if not a > b:
        _fail(f"{a!r} is not greater than {b!r}", reason)


def gt_eq(a: Any, b: Any, reason: Optional[str] = None) -> None:
    if not a >= b:
        _fail(f"{a!r} is not greater than or equal to {b!r}", reason)


def lt(a: Any, b: Any, reason: Optional[str] = None) -> None:
    if not a < b:
        _fail(f"{a!r} is not less than {b!r}", reason)


def lt_eq(a: Any, b: Any, reason: Optional[str] = None) -> None:
    if not a <= b:
        _fail(f"{a!r} is not less than or equal to {b!r}", reason)


def between(val: Any, lo: Any, hi: Any, reason: Optional[str] = None) -> None:
    if not lo <= val <= hi:
        _fail(f"{val!r} is not between {lo!r} and {hi!r}", reason)


def is_in(val: Any, container: Container, reason: Optional[str] = None) -> None:
    if val not in container:
        _fail(f"{val!r} is not in {container!r}", reason)


def is_not_in(val: Any, container: Container, reason: Optional[str] = None) -> None:
    if val in container:
        _fail(f"{val!r} is in {container!r}", reason)


def is_instance(val: Any, types: Union[Type, Tuple[Type, ...]], reason: Optional[str] = None) -> None:
    if not isinstance(val, types):
        _fail(f"{val!r} is not an instance of {types!r}", reason)


def is_subclass(val: Any, parent: Type, reason: Optional[str] = None) -> None:
    if not (isinstance(val, type) and issubclass(val, parent)):
        _fail(f"{val!r} is not a subclass of {parent!r}", reason)


def is_type(val: Any, t: Type, reason: Optional[str] = None) -> None:
    if type(val) is not t:
        _fail(f"{val!r} is not of type {t!r}", reason)


def len_eq(val: Sized, n: int, reason: Optional[str] = None) -> None:
    if len(val) != n:
        _fail(f"length of {val!r} is {len(val)}, not {n}", reason)


def is_empty(val: Sized, reason: Optional[str] = None) -> None:
    if len(val) != 0:
        _fail(f"{val!r} is not empty", reason)


def is_not_empty(val: Sized, reason: Optional[str] = None) -> None:
    if len(val) == 0:
        _fail("collection is empty", reason)


def all_unique(vals: Iterable, reason: Optional[str] = None) -> None:
    seen = list(vals)
    if len(seen) != len(set(seen)):
        _fail(f"{seen!r} contains duplicates", reason)


def check_true(val: Any, reason: Optional[str] = None) -> None:  # legacy alias used by reference call sites
    true(val, reason)
