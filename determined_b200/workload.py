"""The workload protocol: the only contract between the master (searcher + sequencer) and a running
trial. A trial consumes a stream of four workload kinds and answers each with a metrics dict.

Parity: reference ``harness/determined/workload.py:9-236`` and ``master/pkg/workload/workload.go:12-20``
(kinds RUN_STEP / COMPUTE_VALIDATION_METRICS / CHECKPOINT_MODEL / TERMINATE, the
``(workload, args, respond)`` stream shape, ``Skipped`` for non-chief ranks).
"""
from __future__ import annotations

import enum
import pathlib
from typing import Any, Callable, Dict, Iterator, List, Optional, Tuple, Union


class Workload:
    @enum.unique
    class Kind(enum.Enum):
        RUN_STEP = 1
        COMPUTE_VALIDATION_METRICS = 2
        CHECKPOINT_MODEL = 3
        TERMINATE = 4

    __slots__ = ("kind", "experiment_id", "trial_id", "step_id", "num_batches", "total_batches_processed")

    def __init__(self, kind: "Workload.Kind", experiment_id: int, trial_id: int, step_id: int, num_batches: int = 0,
                 total_batches_processed: int = 0) -> None:
        self.kind = kind
        self.experiment_id = experiment_id
        self.trial_id = trial_id
        self.step_id = step_id
        self.num_batches = num_batches
        self.total_batches_processed = total_batches_processed

    def _key(self) -> Tuple:
        return tuple(getattr(self, s) for s in self.__slots__)

    def __eq__(self, other: object) -> bool:
        return isinstance(other, Workload) and self._key() == other._key()

    def __hash__(self) -> int:
        return hash((self.kind, self.experiment_id, self.trial_id, self.step_id))

    def __repr__(self) -> str:
        extra = f" ({self.num_batches} batches)" if self.kind is Workload.Kind.RUN_STEP else ""
        return f"<{self.kind.name}{extra}: ({self.experiment_id},{self.trial_id},{self.step_id})>"

    def to_json(self) -> Dict[str, Any]:
        d = {s: getattr(self, s) for s in self.__slots__}
        d["kind"] = self.kind.name
        return d

    __json__ = to_json

    @staticmethod
    def from_json(d: Dict[str, Any]) -> "Workload":
        if d["kind"] not in Workload.Kind.__members__:
            raise ValueError(f"unknown workload kind {d['kind']!r}")
        return Workload(Workload.Kind[d["kind"]], d["experiment_id"], d["trial_id"], d["step_id"],
                        d.get("num_batches", 0), d.get("total_batches_processed", 0))


Metrics = Dict[str, Any]


class Skipped:
    """Response of a rank that is not the chief: the launcher keeps the single non-Skipped reply."""

    def __repr__(self) -> str:
        return "Skipped()"


Response = Union[Metrics, Skipped]
ResponseFunc = Callable[[Response], None]
Args = List[Any]
Stream = Iterator[Tuple[Workload, Args, ResponseFunc]]


class Source:
    """A harness layer: iterating it yields ``(workload, args, respond)`` for the layer below."""

    def __iter__(self) -> Stream:  # pragma: no cover - interface
        raise NotImplementedError


def train_workload(step_id: int, exp_id: int = 1, trial_id: int = 1, num_batches: int = 1,
                   total_batches_processed: int = 0) -> Workload:
    return Workload(Workload.Kind.RUN_STEP, exp_id, trial_id, step_id, num_batches, total_batches_processed)


def validation_workload(step_id: int = 1, exp_id: int = 1, trial_id: int = 1, total_batches_processed: int = 0) -> Workload:
    return Workload(Workload.Kind.COMPUTE_VALIDATION_METRICS, exp_id, trial_id, step_id, 0, total_batches_processed)


def checkpoint_workload(step_id: int = 1, exp_id: int = 1, trial_id: int = 1, total_batches_processed: int = 0) -> Workload:
    return Workload(Workload.Kind.CHECKPOINT_MODEL, exp_id, trial_id, step_id, 0, total_batches_processed)


def terminate_workload(step_id: int = 1, exp_id: int = 1, trial_id: int = 1) -> Workload:
    return Workload(Workload.Kind.TERMINATE, exp_id, trial_id, step_id)


def ignore_workload_response(_: Response) -> None:
    return None


class WorkloadResponseInterceptor:
    """Test/driver helper: send one workload, remember the answer.

    ``for wkld, args, respond in interceptor.send(train_workload(1)): yield wkld, args, respond``
    then ``interceptor.metrics_result()``. (Reference: ``workload.py:119-185``.)
    """

    def __init__(self) -> None:
        self._response: Optional[Response] = None

    def _respond(self, r: Response) -> None:
        if self._response is not None:
            raise RuntimeError("workload answered twice")
        self._response = r

    def send(self, workload: Workload, args: Optional[Args] = None) -> Stream:
        self._response = None
        yield workload, list(args or []), self._respond
        if self._response is None:
            raise RuntimeError(f"workload {workload} was never answered")

    def metrics_result(self) -> Metrics:
        if self._response is None:
            raise RuntimeError("no response recorded")
        if isinstance(self._response, Skipped):
            raise RuntimeError("response was Skipped")
        return self._response

    def result(self) -> Response:
        if self._response is None:
            raise RuntimeError("no response recorded")
        return self._response


def make_checkpoint_args(path: Union[str, pathlib.Path]) -> Args:
    return [pathlib.Path(path)]
