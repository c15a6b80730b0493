"""The master must survive what arrives on its socket: malformed HTTP, hostile header values, JSON of the wrong shape on the
routes that take bodies, and valid experiment configs with fields deleted or replaced by nonsense. After the barrage it still
answers, and rejects garbage with 4xx instead of dying. (The same generators ran 5,500 requests against an
AddressSanitizer + UBSan build without a report; here a few hundred run against the regular binary on every test run.)"""
import copy
import json
import os
import pathlib
import random
import socket
import subprocess
import sys
import time

import yaml

from determined_b200.layers._master_client import MasterClient

ROOT = pathlib.Path(__file__).resolve().parent.parent
LIB = ROOT / "determined_b200" / "_lib"


def _send(port: int, raw: bytes) -> bytes:
    try:
        s = socket.create_connection(("127.0.0.1", port), timeout=1)
        s.sendall(raw)
        try:
            out = s.recv(4096)
        except OSError:
            out = b""
        s.close()
        return out
    except OSError:
        return b""


def test_master_survives_malformed_requests_and_mutated_configs(tmp_path):
    from determined_b200 import _build

    _build.build_core()
    port_file = tmp_path / "port"
    proc = subprocess.Popen([str(LIB / "det-master"), "--port", "0", "--data-dir", str(tmp_path / "data"), "--slot-type", "none",
                             "--repo-root", str(ROOT), "--python", sys.executable, "--print-port-file", str(port_file)],
                            stdout=open(tmp_path / "master.log", "w"), stderr=subprocess.STDOUT,
                            env=dict(os.environ, DET_CONFIG_DIR=str(tmp_path / "cfg")))
    try:
        for _ in range(200):
            if port_file.exists() and port_file.read_text().strip():
                break
            time.sleep(0.05)
        port = int(port_file.read_text())
        rng = random.Random(20261107)
        paths = ["/api/v1/experiments", "/api/v1/experiments/1", "/api/v1/trials/1/logs?limit=-5&offset=99999999999999999999", "/api/v1/agents",
                 "/api/v1/trials/abc/next_workload", "/proxy/zzz/", "/api/v1/users/login", "/api/v1/checkpoints/../../etc/passwd",
                 "/api/v1/experiments/18446744073709551616", "/api/v1/agents/a/events", "/api/v1/trials/1/workload_completed", "/debug/actors"]
        bodies = [b"", b"{", b"[]", b"null", b'{"a":' + b"[" * 3000, b'"\\ud800"', b"\xff\xfe\x00", b'{"wait_ms": -1, "ttl_ms": 1e99}',
                  b'{"events":[{"container_id":5}],"logs":"x"}', b'{"metrics": {"a": NaN}}', b"1e99999", b'{"k":"' + b"\\" * 999 + b'"}']
        for _ in range(80):                       # wire level
            m, p, b = rng.choice(["GET", "POST", "PUT", "DELETE", "BREW", ""]), rng.choice(paths), rng.choice(bodies)
            kind = rng.random()
            if kind < 0.55:
                raw = f"{m} {p} HTTP/1.1\r\nHost: x\r\nContent-Length: {len(b)}\r\n\r\n".encode() + b
            elif kind < 0.7:
                raw = f"{m} {p} HTTP/1.1\r\nContent-Length: {rng.choice(['-1', '99999999999999999999', 'abc'])}\r\n\r\n".encode() + b[:3]
            elif kind < 0.85:
                raw = bytes(rng.getrandbits(8) for _ in range(rng.randint(1, 300)))
            else:
                raw = f"{m} {p} HTTP/1.1\r\n".encode() + b"X: " + b"a" * rng.randint(1, 70000) + b"\r\n\r\n"
            _send(port, raw)
        base = yaml.safe_load((ROOT / "tests" / "fixtures" / "no_op" / "adaptive_chaos.yaml").read_text())
        junk = [None, -1, 0, 2 ** 40, 1e-300, "", "pbt", "grid", [], {}, {"batches": -5}, {"epochs": 0}, {"type": "int", "minval": 9, "maxval": 1},
                {"type": "log", "minval": 5, "maxval": -5, "base": 0}, {"type": "categorical", "vals": []}, True, "x" * 300, {"records": 2 ** 62}]
        statuses = []
        for _ in range(100):                      # config level: one to three fields deleted or replaced
            cfg = copy.deepcopy(base)
            for _ in range(rng.randint(1, 3)):
                node = cfg
                while isinstance(node, dict) and node:
                    k = rng.choice(list(node))
                    if not isinstance(node[k], dict) or rng.random() < 0.35:
                        if rng.random() < 0.2:
                            del node[k]
                        else:
                            node[k] = rng.choice(junk)
                        break
                    node = node[k]
            route = rng.choice(["/api/v1/preview_search", "/api/v1/experiments"])
            body = json.dumps({"config": cfg, "model_definition": [{"path": "model_def.py", "content": "x = 1\n"}], "activate": False}).encode()
            out = _send(port, f"POST {route} HTTP/1.1\r\nHost: x\r\nContent-Length: {len(body)}\r\n\r\n".encode() + body)
            if out.startswith(b"HTTP/1.1 "):
                statuses.append(int(out[9:12]))
        assert proc.poll() is None, (tmp_path / "master.log").read_text()[-2000:]
        assert statuses and all(s in (200, 201, 400, 404, 409, 413, 422) for s in statuses), sorted(set(statuses))
        assert any(s == 400 for s in statuses), "nonsense configs must be rejected, not accepted"
        assert MasterClient("127.0.0.1", port).request("GET", "/info")      # still serving
    finally:
        proc.terminate()
        try:
            proc.wait(timeout=5)
        except Exception:
            proc.kill()
