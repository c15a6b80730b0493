"""TLS on the control plane: the master serves HTTPS from ``security.tls.{cert,key}``; SDK / CLI sessions, the harness's
workload channel and det-agent pin its certificate (reference: master/internal/config.go:252-256,
agent/internal/options.go:12-103, common/determined_common/api/request.py:93-215)."""
import datetime
import ipaddress
import os
import pathlib
import subprocess
import sys
import time

import pytest

from determined_b200.experimental.client import APIError, Determined

ROOT = pathlib.Path(__file__).resolve().parent.parent
LIB = ROOT / "determined_b200" / "_lib"
MODEL_DIR = str(ROOT / "tests" / "fixtures" / "xor_model_def")


def make_cert(dirpath: pathlib.Path, name: str):
    from cryptography import x509
    from cryptography.hazmat.primitives import hashes, serialization
    from cryptography.hazmat.primitives.asymmetric import ec
    from cryptography.x509.oid import NameOID

    key = ec.generate_private_key(ec.SECP256R1())
    subject = x509.Name([x509.NameAttribute(NameOID.COMMON_NAME, name)])
    now = datetime.datetime.now(datetime.timezone.utc)
    cert = (x509.CertificateBuilder().subject_name(subject).issuer_name(subject).public_key(key.public_key())
            .serial_number(x509.random_serial_number()).not_valid_before(now - datetime.timedelta(minutes=5))
            .not_valid_after(now + datetime.timedelta(days=2))
            .add_extension(x509.SubjectAlternativeName([x509.DNSName(name), x509.IPAddress(ipaddress.ip_address("127.0.0.1"))]), critical=False)
            .add_extension(x509.BasicConstraints(ca=True, path_length=None), critical=True)
            .sign(key, hashes.SHA256()))
    c, k = dirpath / f"{name}.crt", dirpath / f"{name}.key"
    c.write_bytes(cert.public_bytes(serialization.Encoding.PEM))
    k.write_bytes(key.private_bytes(serialization.Encoding.PEM, serialization.PrivateFormat.PKCS8, serialization.NoEncryption()))
    return str(c), str(k)


@pytest.fixture(scope="module")
def tls_master(tmp_path_factory):
    from determined_b200 import _build

    _build.build_core()
    tmp = tmp_path_factory.mktemp("tlsmaster")
    cert, key = make_cert(tmp, "det-master.test")
    other_cert, _ = make_cert(tmp, "someone-else.test")
    port_file = tmp / "port"
    proc = subprocess.Popen([str(LIB / "det-master"), "--port", "0", "--data-dir", str(tmp / "data"), "--artificial-slots", "2",
                             "--repo-root", str(ROOT), "--python", sys.executable, "--print-port-file", str(port_file),
                             "--security-tls-cert", cert, "--security-tls-key", key],
                            stdout=open(tmp / "master.log", "w"), stderr=subprocess.STDOUT, env=dict(os.environ, DET_CONFIG_DIR=str(tmp / "cfg")))
    for _ in range(100):
        if port_file.exists() and port_file.read_text().strip():
            break
        time.sleep(0.05)
    port = int(port_file.read_text())
    old = {k: os.environ.get(k) for k in ("DET_CONFIG_DIR", "DET_MASTER_CERT_FILE", "DET_MASTER_CERT_NAME")}
    os.environ["DET_CONFIG_DIR"] = str(tmp / "cfg")
    yield {"port": port, "cert": cert, "other_cert": other_cert, "tmp": tmp}
    proc.terminate()
    try:
        proc.wait(timeout=5)
    except Exception:
        proc.kill()
    for k, v in old.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = v


def test_https_session_pins_the_master_certificate(tls_master, monkeypatch):
    port = tls_master["port"]
    monkeypatch.setenv("DET_MASTER_CERT_FILE", tls_master["cert"])
    d = Determined(f"https://127.0.0.1:{port}")
    assert d._session.get("/api/v1/master")["cluster_name"]
    # the certificate also names det-master.test: dial the address, expect that name
    monkeypatch.setenv("DET_MASTER_CERT_NAME", "det-master.test")
    assert Determined(f"https://127.0.0.1:{port}")._session.get("/api/v1/master")["version"]
    # a name the certificate does not carry is refused
    monkeypatch.setenv("DET_MASTER_CERT_NAME", "not-the-master.test")
    with pytest.raises((APIError, OSError, Exception)) as e:
        Determined(f"https://127.0.0.1:{port}")._session.get("/api/v1/master")
    assert "certificate" in str(e.value).lower() or "hostname" in str(e.value).lower() or "ssl" in str(e.value).lower()
    monkeypatch.delenv("DET_MASTER_CERT_NAME")
    # another CA: refused; noverify: accepted (encryption only); plain HTTP to the TLS port: no answer
    monkeypatch.setenv("DET_MASTER_CERT_FILE", tls_master["other_cert"])
    with pytest.raises(Exception):
        Determined(f"https://127.0.0.1:{port}")._session.get("/api/v1/master")
    monkeypatch.setenv("DET_MASTER_CERT_FILE", "noverify")
    assert Determined(f"https://127.0.0.1:{port}")._session.get("/api/v1/master")["version"]
    with pytest.raises(Exception):
        Determined(f"http://127.0.0.1:{port}")._session.request("GET", "/api/v1/master", timeout=5)


def test_experiment_runs_with_the_harness_on_tls(tls_master, monkeypatch):
    monkeypatch.setenv("DET_MASTER_CERT_FILE", tls_master["cert"])
    d = Determined(f"https://127.0.0.1:{tls_master['port']}")
    cfg = {"entrypoint": "model_def:XorTrial", "description": "tls", "scheduling_unit": 10, "max_restarts": 0,
           "hyperparameters": {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5},
           "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 20}},
           "checkpoint_storage": {"type": "shared_fs", "host_path": str(tls_master["tmp"] / "ckpts")}}
    exp = d.create_experiment(cfg, MODEL_DIR)
    assert exp.wait(timeout=180) == "COMPLETED"
    t = exp.describe()["trials"][0]
    assert t["total_batches_processed"] == 20 and t["restarts"] == 0


def test_agent_registers_over_tls_and_is_refused_with_the_wrong_certificate(tls_master, monkeypatch):
    monkeypatch.setenv("DET_MASTER_CERT_FILE", tls_master["cert"])
    d = Determined(f"https://127.0.0.1:{tls_master['port']}")
    common = [str(LIB / "det-agent"), "--master-host", "127.0.0.1", "--master-port", str(tls_master["port"]), "--artificial-slots", "1",
              "--slot-type", "none", "--security-tls-enabled"]
    good = subprocess.Popen(common + ["--agent-id", "agent-tls", "--security-tls-master-cert", tls_master["cert"]],
                            stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    bad = subprocess.Popen(common + ["--agent-id", "agent-badca", "--security-tls-master-cert", tls_master["other_cert"]],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    try:
        ids = []
        for _ in range(60):
            ids = [a["id"] for a in d._session.get("/api/v1/agents")]
            if "agent-tls" in ids:
                break
            time.sleep(0.25)
        assert "agent-tls" in ids and "agent-badca" not in ids
    finally:
        good.terminate()
        bad.terminate()
        out_bad = bad.communicate(timeout=10)[0]
        good.communicate(timeout=10)
    assert "certificate verification failed" in out_bad


def test_auth_required_exempts_only_task_and_agent_callbacks(tmp_path):
    """With ``auth_required`` every route needs a session except login, the public info pages, the harness's workload
    exchange / log shipping and the agents' register / commands / events calls. State-changing routes under the same path
    prefixes (trial kill, agent unregister, slot enable / disable, master config, tracing) are not exempt."""
    import json
    import urllib.error
    import urllib.request

    port_file = tmp_path / "port"
    proc = subprocess.Popen([str(LIB / "det-master"), "--port", "0", "--data-dir", str(tmp_path / "data"), "--artificial-slots", "1",
                             "--repo-root", str(ROOT), "--python", sys.executable, "--print-port-file", str(port_file), "--auth-required"],
                            stdout=open(tmp_path / "master.log", "w"), stderr=subprocess.STDOUT)
    try:
        for _ in range(100):
            if port_file.exists() and port_file.read_text().strip():
                break
            time.sleep(0.05)
        base = f"http://127.0.0.1:{int(port_file.read_text())}"

        def call(method, path, body=None, token=None):
            req = urllib.request.Request(base + path, method=method, data=None if body is None else json.dumps(body).encode(),
                                         headers=dict({"Content-Type": "application/json"}, **({"Authorization": f"Bearer {token}"} if token else {})))
            try:
                with urllib.request.urlopen(req, timeout=10) as r:
                    return r.status
            except urllib.error.HTTPError as e:
                return e.code

        assert call("GET", "/info") == 200 and call("GET", "/api/v1/master") == 200
        for method, path in (("GET", "/api/v1/master/config"), ("POST", "/api/v1/trials/1/kill"), ("POST", "/api/v1/agents/agent-local/unregister"),
                             ("POST", "/api/v1/agents/agent-local/slots/0/disable"), ("GET", "/api/v1/agents"), ("POST", "/debug/trace"),
                             ("GET", "/api/v1/experiments")):
            assert call(method, path, {} if method == "POST" else None) == 401, (method, path)
        # task / agent callbacks are reachable without a session (they answer with their own errors, never 401)
        assert call("POST", "/api/v1/trial_logs", {"logs": []}) != 401
        assert call("GET", "/api/v1/trials/1/next_workload?container_id=x") != 401
        assert call("POST", "/api/v1/agents/agent-x/register", {"devices": [], "label": "", "addresses": ["127.0.0.1"]}) != 401
        req = urllib.request.Request(base + "/api/v1/auth/login", method="POST", data=json.dumps({"username": "determined", "password": ""}).encode())
        token = json.loads(urllib.request.urlopen(req, timeout=10).read())["token"]
        assert call("GET", "/api/v1/master/config", token=token) == 200 and call("GET", "/api/v1/agents", token=token) == 200
    finally:
        proc.terminate()
        try:
            proc.wait(timeout=5)
        except Exception:
            proc.kill()
