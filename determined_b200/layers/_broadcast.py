"""Fan a workload stream out to every worker process of a trial.

The reference does this with a launcher process and ZMQ PUB/SUB + PUSH/PULL sockets between it and the
``horovodrun``-spawned workers (``layers/_worker_process.py:75-339``, ``ipc.py:63-222``), keeping exactly one
non-``Skipped`` reply. Here every rank runs the same loop: rank 0 owns the real stream (master channel + workload
manager) and broadcasts each ``(workload, args)`` over the trial's control plane; the other ranks iterate the
broadcasts and their (``Skipped``) replies are dropped locally. A ``None`` broadcast ends the stream; an exception on
the chief is broadcast so that peers fail fast instead of hanging in a collective.
"""
from __future__ import annotations

from typing import Any, Optional

from .. import workload as wl


class _Failure:
    def __init__(self, text: str) -> None:
        self.text = text


class BroadcastWorkloads(wl.Source):
    def __init__(self, comm: Any, chief_stream: Optional[wl.Stream]) -> None:
        self.comm = comm
        self.chief_stream = chief_stream

    def __iter__(self) -> wl.Stream:
        if self.comm.size <= 1:
            assert self.chief_stream is not None
            yield from self.chief_stream
            return
        if self.comm.rank == 0:
            assert self.chief_stream is not None
            try:
                for workload, args, respond in self.chief_stream:
                    self.comm.broadcast_object((workload.to_json(), [str(a) for a in args]), 0)
                    yield workload, args, respond
            except BaseException as e:
                try:
                    self.comm.broadcast_object(_Failure(repr(e)), 0)
                finally:
                    raise
            self.comm.broadcast_object(None, 0)
        else:
            import pathlib

            while True:
                item = self.comm.broadcast_object(None, 0)
                if item is None:
                    return
                if isinstance(item, _Failure):
                    raise RuntimeError(f"chief worker failed: {item.text}")
                wjson, args = item
                workload = wl.Workload.from_json(wjson)
                yield workload, [pathlib.Path(a) for a in args], lambda _r: None
                if workload.kind == wl.Workload.Kind.TERMINATE:
                    return
