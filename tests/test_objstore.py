"""Checkpoint storage on S3 / GCS / HDFS through the stdlib clients of ``utils/objstore.py``, against in-process stand-in
servers (the S3 one verifies every SigV4 signature independently)."""
import os

import pytest

from determined_b200 import storage
from determined_b200.utils import objstore
from tests import fake_objstore


def _round_trip(mgr: storage.StorageManager, tmp_path) -> storage.StorageMetadata:
    with mgr.store_path() as (sid, path):
        os.makedirs(os.path.join(path, "sub/empty"))
        open(os.path.join(path, "state_dict.pth"), "wb").write(b"\x00\x01" * 5000)
        open(os.path.join(path, "sub", "näme with space.txt"), "w").write("hello")
    resources = {"state_dict.pth": 10000, "sub/": 0, "sub/empty/": 0, "sub/näme with space.txt": 5}
    meta = storage.StorageMetadata(sid, resources)
    with mgr.restore_path(meta) as path:
        assert open(os.path.join(path, "state_dict.pth"), "rb").read() == b"\x00\x01" * 5000
        assert open(os.path.join(path, "sub", "näme with space.txt")).read() == "hello"
        assert os.path.isdir(os.path.join(path, "sub", "empty"))
    return meta


def test_s3_checkpoint_round_trip_signed(tmp_path):
    srv, url, objects, stats = fake_objstore.make_s3()
    try:
        mgr = storage.build({"type": "s3", "bucket": "ckpts", "access_key": "AKIDEXAMPLE", "secret_key": "s3cr3t",
                             "endpoint_url": url, "prefix": "team-a", "save_trial_best": 1})
        meta = _round_trip(mgr, tmp_path)
        assert stats["bad_signature"] == 0 and stats["requests"] >= 6
        assert f"team-a/{meta.storage_id}/state_dict.pth" in objects
        assert sorted(k for k, _ in mgr.client.list(f"team-a/{meta.storage_id}/")) == sorted(
            f"team-a/{meta.storage_id}/{r}" for r in meta.resources)            # pages of 2 → continuation tokens
        assert mgr.client.exists(f"team-a/{meta.storage_id}/state_dict.pth") and not mgr.client.exists("nope")
        mgr.delete(meta)
        assert not objects
        storage.validate_config({"type": "s3", "bucket": "ckpts", "access_key": "AKIDEXAMPLE", "secret_key": "s3cr3t", "endpoint_url": url})
        assert not objects
    finally:
        srv.shutdown()


def test_s3_wrong_secret_is_rejected_and_multipart(tmp_path, monkeypatch):
    srv, url, objects, stats = fake_objstore.make_s3()
    try:
        bad = objstore.S3Client("b", "AKIDEXAMPLE", "wrong", url)
        with pytest.raises(objstore.ObjectStoreError) as e:
            bad.put_bytes("k", b"x")
        assert e.value.status == 403 and stats["bad_signature"] == 1
        monkeypatch.setattr(objstore, "MULTIPART_THRESHOLD", 1 << 16)
        monkeypatch.setattr(objstore, "PART_SIZE", 1 << 16)
        blob = os.urandom(3 * (1 << 16) + 123)
        src = tmp_path / "big.bin"
        src.write_bytes(blob)
        c = objstore.S3Client("b", "AKIDEXAMPLE", "s3cr3t", url)
        c.put_file("dir/big.bin", str(src))
        assert stats["parts"] == 4 and objects["dir/big.bin"] == blob
        c.get_file("dir/big.bin", str(tmp_path / "back.bin"))
        assert (tmp_path / "back.bin").read_bytes() == blob
    finally:
        srv.shutdown()


def test_gcs_checkpoint_round_trip(tmp_path, monkeypatch):
    srv, url, objects = fake_objstore.make_gcs()
    monkeypatch.setenv("STORAGE_EMULATOR_HOST", url)
    try:
        mgr = storage.build({"type": "gcs", "bucket": "ckpts"})
        meta = _round_trip(mgr, tmp_path)
        assert len(list(mgr.client.list(meta.storage_id))) == 4
        mgr.delete(meta)
        assert not objects
    finally:
        srv.shutdown()


def test_gcs_service_account_grant(tmp_path, monkeypatch):
    """The JWT grant: a key file → RS256 assertion posted to token_uri → bearer token on the request."""
    import json
    import threading
    from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

    from cryptography.hazmat.primitives import serialization
    from cryptography.hazmat.primitives.asymmetric import rsa

    seen = {}

    class Tok(BaseHTTPRequestHandler):
        def log_message(self, *a):
            pass

        def do_POST(self):
            seen["form"] = self.rfile.read(int(self.headers["Content-Length"])).decode()
            out = json.dumps({"access_token": "tok-123", "expires_in": 3600}).encode()
            self.send_response(200)
            self.send_header("Content-Length", str(len(out)))
            self.end_headers()
            self.wfile.write(out)

    tsrv = ThreadingHTTPServer(("127.0.0.1", 0), Tok)
    threading.Thread(target=tsrv.serve_forever, daemon=True).start()
    key = rsa.generate_private_key(public_exponent=65537, key_size=2048)
    pem = key.private_bytes(serialization.Encoding.PEM, serialization.PrivateFormat.PKCS8, serialization.NoEncryption()).decode()
    kf = tmp_path / "sa.json"
    kf.write_text(json.dumps({"client_email": "svc@proj.iam", "private_key": pem,
                              "token_uri": f"http://127.0.0.1:{tsrv.server_address[1]}/token"}))
    monkeypatch.delenv("STORAGE_EMULATOR_HOST", raising=False)
    monkeypatch.delenv("GOOGLE_OAUTH_ACCESS_TOKEN", raising=False)
    monkeypatch.setenv("GOOGLE_APPLICATION_CREDENTIALS", str(kf))
    try:
        c = objstore.GCSClient("b")
        assert c._bearer() == "tok-123" and "jwt-bearer" in seen["form"] and "assertion=" in seen["form"]
    finally:
        tsrv.shutdown()


def test_hdfs_checkpoint_round_trip(tmp_path):
    srv, url, files, stats = fake_objstore.make_webhdfs()
    try:
        mgr = storage.build({"type": "hdfs", "hdfs_url": url, "hdfs_path": "/determined/ckpts", "user": "det"})
        meta = _round_trip(mgr, tmp_path)
        assert f"/determined/ckpts/{meta.storage_id}/state_dict.pth" in files and stats["redirects"] >= 4
        mgr.delete(meta)
        assert not files
    finally:
        srv.shutdown()
