"""TLS for the Python clients of the master (SDK / CLI session, the harness's workload channel).

Reference behaviour (``common/determined_common/api/request.py:93-215``, ``harness/determined/layers/_socket_manager.py``):
the master URL may be ``https://``; the certificate to trust is pinned — a file handed to the task by the master
(``DET_MASTER_CERT_FILE``) or kept by the CLI — and the name to expect in it may differ from the address being dialled
(``DET_MASTER_CERT_NAME``: tasks reach the master by IP while its certificate names a host). ``noverify`` as the file
switches verification off and only encrypts.
"""
from __future__ import annotations

import http.client
import os
import ssl
import urllib.request
from typing import Any, Optional


def make_context(cert_file: Optional[str]) -> ssl.SSLContext:
    if cert_file == "noverify":
        ctx = ssl.create_default_context()
        ctx.check_hostname = False
        ctx.verify_mode = ssl.CERT_NONE
        return ctx
    return ssl.create_default_context(cafile=cert_file or None)


class _NamedHTTPSConnection(http.client.HTTPSConnection):
    """HTTPS connection whose certificate is checked against ``cert_name`` instead of the host in the URL."""

    def __init__(self, *args: Any, cert_name: Optional[str] = None, **kwargs: Any) -> None:
        super().__init__(*args, **kwargs)
        self._cert_name = cert_name

    def connect(self) -> None:
        http.client.HTTPConnection.connect(self)
        self.sock = self._context.wrap_socket(self.sock, server_hostname=self._cert_name or self.host)  # type: ignore[attr-defined]


class _NamedHTTPSHandler(urllib.request.HTTPSHandler):
    def __init__(self, context: ssl.SSLContext, cert_name: Optional[str]) -> None:
        super().__init__(context=context)
        self._ctx, self._cert_name = context, cert_name

    def https_open(self, req: urllib.request.Request) -> Any:
        def conn(host: str, **kw: Any) -> http.client.HTTPSConnection:
            kw.pop("context", None)
            return _NamedHTTPSConnection(host, cert_name=self._cert_name, context=self._ctx, **kw)

        return self.do_open(conn, req)


def build_opener(cert_file: Optional[str] = None, cert_name: Optional[str] = None) -> urllib.request.OpenerDirector:
    """An opener for ``http://`` and ``https://`` master URLs. Arguments default to ``DET_MASTER_CERT_FILE`` /
    ``DET_MASTER_CERT_NAME``."""
    cert_file = cert_file if cert_file is not None else os.environ.get("DET_MASTER_CERT_FILE")
    cert_name = cert_name if cert_name is not None else os.environ.get("DET_MASTER_CERT_NAME")
    return urllib.request.build_opener(_NamedHTTPSHandler(make_context(cert_file), cert_name or None))
