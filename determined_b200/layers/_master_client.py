"""Harness ↔ master channel: a workload ``Source`` that long-polls the master for the next workload and posts the
``WORKLOAD_COMPLETED`` message back.

Role of the reference's ``SocketManager`` (``harness/determined/layers/_socket_manager.py:34-207``: WebSocket to
``/ws/trial/{exp}/{trial}/{container}``, ``RUN_WORKLOAD`` in, ``WORKLOAD_COMPLETED`` JSON out). On one box the
WebSocket is replaced by two HTTP endpoints (``GET .../next_workload`` long-poll, ``POST .../workload_completed``);
the message bodies keep the reference's shape.
"""
from __future__ import annotations

import datetime
import json
import time
import urllib.error
import urllib.request
from typing import Any, Dict, Optional

from .. import workload as wl
from .._env import EnvContext
from ..utils import util


def _now() -> str:
    return datetime.datetime.now(datetime.timezone.utc).isoformat()


class MasterClient:
    def __init__(self, host: str, port: int, timeout: float = 40.0, use_tls: bool = False, cert_file: Optional[str] = None,
                 cert_name: Optional[str] = None) -> None:
        from ..utils import tls

        self.base = f"{'https' if use_tls else 'http'}://{host}:{port}"
        self.timeout = timeout
        self._opener = tls.build_opener(cert_file, cert_name)

    def request(self, method: str, path: str, body: Any = None, retries: int = 5) -> Any:
        data = None if body is None else util.json_encode(body).encode()
        last: Optional[Exception] = None
        for attempt in range(retries):
            try:
                req = urllib.request.Request(self.base + path, data=data, method=method,
                                             headers={"Content-Type": "application/json"})
                with self._opener.open(req, timeout=self.timeout) as resp:
                    text = resp.read().decode()
                    return json.loads(text) if text else None
            except urllib.error.HTTPError as e:
                raise RuntimeError(f"master answered {e.code} for {method} {path}: {e.read().decode(errors='replace')}") from e
            except Exception as e:  # connection refused / timeout: the master may be restarting
                last = e
                time.sleep(min(0.5 * (attempt + 1), 3.0))
        raise RuntimeError(f"master unreachable for {method} {path}: {last!r}")


class MasterWorkloadSource(wl.Source):
    """Yields ``(workload, [], respond)`` until the master says TERMINATE."""

    def __init__(self, env: EnvContext, client: Optional[MasterClient] = None) -> None:
        self.env = env
        self.client = client or MasterClient(env.master_addr, env.master_port, use_tls=bool(getattr(env, "use_tls", False)),
                                             cert_file=getattr(env, "master_cert_file", None),
                                             cert_name=getattr(env, "master_cert_name", None))
        self.trial_id = env.det_trial_id

    current: Optional[wl.Workload] = None

    def report_invalid_hp(self) -> None:
        """The trial raised ``det.InvalidHP``: answer the in-flight workload with ``exited_reason = INVALID_HP`` so the
        searcher can move on (the sequencer then checkpoints nothing and closes the trial gracefully)."""
        if self.current is None:
            return
        self.client.request("POST", f"/api/v1/trials/{self.trial_id}/workload_completed",
                            {"type": "WORKLOAD_COMPLETED", "workload": self.current.to_json(), "exited_reason": "INVALID_HP",
                             "metrics": {}, "start_time": _now(), "end_time": _now()})

    def __iter__(self) -> wl.Stream:
        while True:
            msg = self.client.request("GET", f"/api/v1/trials/{self.trial_id}/next_workload?container_id={self.env.container_id}")
            kind = (msg or {}).get("type", "WAIT")
            if kind == "WAIT":
                continue
            if kind == "TERMINATE":
                w = msg.get("workload")
                workload = wl.Workload.from_json(w) if w else wl.terminate_workload()
                yield workload, [], lambda _r: None
                return
            if kind != "RUN_WORKLOAD":
                raise RuntimeError(f"unexpected message from master: {msg}")
            workload = wl.Workload.from_json(msg["workload"])
            self.current = workload
            start = _now()
            box: Dict[str, Any] = {}

            def respond(response: wl.Response, _w=workload, _start=start) -> None:
                if isinstance(response, wl.Skipped):
                    return
                body = dict(response)
                body.setdefault("type", "WORKLOAD_COMPLETED")
                body["workload"] = _w.to_json()
                body.setdefault("start_time", _start)
                body.setdefault("end_time", _now())
                box["sent"] = True
                self.client.request("POST", f"/api/v1/trials/{self.trial_id}/workload_completed", body)

            yield workload, [], respond
            if not box.get("sent"):
                raise RuntimeError(f"workload {workload} was not answered by the layers below")
