"""``det tunnel`` — a byte tunnel to a service task through the master.

Reference: ``cli/determined_cli/tunnel.py:1-121`` (used as the ssh ``ProxyCommand`` of ``det shell open``, or with
``--listener PORT`` to expose the service on a local port). The master's ``/proxy/<task id>/`` route answers an
``Upgrade: det-tcp`` request with ``101`` after connecting to the task's port, and from then on copies bytes both ways, so
the task can speak any protocol. ``https://`` masters are dialled over TLS with the pinned certificate
(``DET_MASTER_CERT_FILE`` / ``DET_MASTER_CERT_NAME``).
"""
from __future__ import annotations

import os
import select
import socket
import sys
import threading
import urllib.parse
from typing import Optional


def open_tunnel(master: str, task_id: str, token: Optional[str] = None) -> socket.socket:
    """Connected socket whose far end is the task's service port."""
    if "://" not in master:
        master = "http://" + master
    u = urllib.parse.urlparse(master)
    host, port = u.hostname or "127.0.0.1", u.port or (443 if u.scheme == "https" else 8080)
    sock = socket.create_connection((host, port), timeout=30)
    if u.scheme == "https":
        from ..utils import tls

        ctx = tls.make_context(os.environ.get("DET_MASTER_CERT_FILE"))
        sock = ctx.wrap_socket(sock, server_hostname=os.environ.get("DET_MASTER_CERT_NAME") or host)
    req = f"GET /proxy/{task_id}/ HTTP/1.1\r\nHost: {host}:{port}\r\nUpgrade: det-tcp\r\nConnection: Upgrade\r\n"
    if token:
        req += f"Authorization: Bearer {token}\r\n"
    sock.sendall((req + "\r\n").encode())
    head = b""
    while b"\r\n\r\n" not in head:
        chunk = sock.recv(4096)
        if not chunk:
            raise ConnectionError("master closed the connection during the tunnel handshake")
        head += chunk
    status_line, _, rest = head.partition(b"\r\n")
    if b" 101 " not in status_line + b" ":
        raise ConnectionError(f"tunnel refused: {status_line.decode(errors='replace')}")
    sock.settimeout(None)
    leftover = head.split(b"\r\n\r\n", 1)[1]
    if leftover:   # bytes the service sent right behind the handshake (an ssh banner)
        sock = _Prefixed(sock, leftover)  # type: ignore[assignment]
    del rest
    return sock


class _Prefixed:
    """Socket wrapper that hands out bytes already read during the handshake first."""

    def __init__(self, sock: socket.socket, prefix: bytes) -> None:
        self._sock, self._prefix = sock, prefix

    def recv(self, n: int) -> bytes:
        if self._prefix:
            out, self._prefix = self._prefix[:n], self._prefix[n:]
            return out
        return self._sock.recv(n)

    def __getattr__(self, name: str):
        return getattr(self._sock, name)


def _copy(src, dst_write, on_close) -> None:
    try:
        while True:
            data = src.recv(65536)
            if not data:
                break
            dst_write(data)
    except OSError:
        pass
    finally:
        on_close()


def stdio_tunnel(master: str, task_id: str, token: Optional[str] = None) -> int:
    """stdin -> task, task -> stdout (what ssh's ProxyCommand wants)."""
    sock = open_tunnel(master, task_id, token)
    done = threading.Event()
    out = sys.stdout.buffer

    def write_out(b: bytes) -> None:
        out.write(b)
        out.flush()

    threading.Thread(target=_copy, args=(sock, write_out, done.set), daemon=True).start()
    fd = sys.stdin.fileno()
    while not done.is_set():
        r, _, _ = select.select([fd], [], [], 0.2)
        if r:
            data = os.read(fd, 65536)
            if not data:
                break
            sock.sendall(data)
    try:
        sock.shutdown(socket.SHUT_WR)
    except OSError:
        pass
    done.wait(timeout=2)
    return 0


def listen_tunnel(master: str, task_id: str, port: int, token: Optional[str] = None, once: bool = False, ready=None) -> int:
    """Accept connections on 127.0.0.1:port and carry each one through its own tunnel."""
    srv = socket.socket()
    srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
    srv.bind(("127.0.0.1", port))
    srv.listen(16)
    if ready is not None:
        ready(srv.getsockname()[1])
    while True:
        conn, _ = srv.accept()
        try:
            far = open_tunnel(master, task_id, token)
        except Exception as e:  # noqa: BLE001
            print(f"det tunnel: {e}", file=sys.stderr)
            conn.close()
            if once:
                return 1
            continue

        def close_both(c=conn, f=far) -> None:
            for s in (c, f):
                try:
                    s.shutdown(socket.SHUT_RDWR)
                except OSError:
                    pass

        t1 = threading.Thread(target=_copy, args=(conn, far.sendall, close_both), daemon=True)
        t2 = threading.Thread(target=_copy, args=(far, conn.sendall, close_both), daemon=True)
        t1.start()
        t2.start()
        if once:
            t1.join()
            t2.join()
            return 0
