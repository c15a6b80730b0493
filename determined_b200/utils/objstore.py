"""Object-store clients on the standard library: S3 (AWS Signature V4, path-style, multipart upload) and GCS (JSON API with
a bearer token). They are what checkpoint storage (``storage.S3StorageManager`` / ``GCSStorageManager``) and the data-layer
cache (``data_layer.type: s3 | gcs``) talk to when the vendor SDKs (boto3, google-cloud-storage) are not installed — the
reference depends on those SDKs (``common/determined_common/storage/s3.py``, ``gcs.py``; yogadl's S3/GCS storage).

Both clients expose the same five calls: ``put_file``, ``get_file``, ``put_bytes``, ``exists``, ``delete``, ``list``.
"""
from __future__ import annotations

import base64
import datetime
import hashlib
import hmac
import http.client
import json
import os
import ssl
import time
import urllib.parse
import xml.etree.ElementTree as ET
from typing import Any, Dict, Iterator, List, Optional, Tuple

_EMPTY_SHA = hashlib.sha256(b"").hexdigest()
MULTIPART_THRESHOLD = 64 << 20
PART_SIZE = 64 << 20


class ObjectStoreError(RuntimeError):
    def __init__(self, status: int, body: str, what: str) -> None:
        super().__init__(f"{what}: HTTP {status}: {body[:300]}")
        self.status = status


def _connect(scheme: str, host: str, timeout: float) -> http.client.HTTPConnection:
    if scheme == "https":
        return http.client.HTTPSConnection(host, timeout=timeout, context=ssl.create_default_context())
    return http.client.HTTPConnection(host, timeout=timeout)


def _quote(s: str, safe: str = "-_.~") -> str:
    return urllib.parse.quote(s, safe=safe)


def _file_sha256(path: str, offset: int = 0, length: Optional[int] = None) -> str:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        f.seek(offset)
        left = length if length is not None else os.path.getsize(path) - offset
        while left > 0:
            chunk = f.read(min(left, 1 << 20))
            if not chunk:
                break
            h.update(chunk)
            left -= len(chunk)
    return h.hexdigest()


class _FileSlice:
    """A readable window of a file, so one part of a multipart upload streams without being read into memory."""

    def __init__(self, path: str, offset: int, length: int) -> None:
        self.f = open(path, "rb")
        self.f.seek(offset)
        self.left = length

    def read(self, n: int = -1) -> bytes:
        if self.left <= 0:
            return b""
        n = self.left if n is None or n < 0 else min(n, self.left)
        data = self.f.read(n)
        self.left -= len(data)
        return data

    def close(self) -> None:
        self.f.close()


class S3Client:
    """Path-style S3 requests signed with AWS Signature Version 4. Works against AWS and S3-compatible endpoints
    (``endpoint_url``). Credentials: arguments, else ``AWS_ACCESS_KEY_ID`` / ``AWS_SECRET_ACCESS_KEY`` /
    ``AWS_SESSION_TOKEN``; region: argument, ``AWS_DEFAULT_REGION``, else ``us-east-1``."""

    def __init__(self, bucket: str, access_key: Optional[str] = None, secret_key: Optional[str] = None,
                 endpoint_url: Optional[str] = None, region: Optional[str] = None, timeout: float = 120.0) -> None:
        self.bucket = bucket
        self.access_key = access_key or os.environ.get("AWS_ACCESS_KEY_ID", "")
        self.secret_key = secret_key or os.environ.get("AWS_SECRET_ACCESS_KEY", "")
        self.token = os.environ.get("AWS_SESSION_TOKEN") if not access_key else None
        self.region = region or os.environ.get("AWS_DEFAULT_REGION") or os.environ.get("AWS_REGION") or "us-east-1"
        url = urllib.parse.urlparse(endpoint_url or f"https://s3.{self.region}.amazonaws.com")
        self.scheme, self.host, self.base = url.scheme or "https", url.netloc, url.path.rstrip("/")
        self.timeout = timeout
        if not self.access_key or not self.secret_key:
            raise RuntimeError("S3 needs credentials: access_key/secret_key in the config or AWS_ACCESS_KEY_ID/AWS_SECRET_ACCESS_KEY")

    # -- signing -------------------------------------------------------------------------------------------------
    def _sign(self, method: str, path: str, query: Dict[str, str], headers: Dict[str, str], payload_sha: str,
              now: Optional[datetime.datetime] = None) -> Dict[str, str]:
        now = now or datetime.datetime.now(datetime.timezone.utc)
        amz_date, day = now.strftime("%Y%m%dT%H%M%SZ"), now.strftime("%Y%m%d")
        h = {k.lower(): " ".join(str(v).split()) for k, v in headers.items()}
        h.update({"host": self.host, "x-amz-date": amz_date, "x-amz-content-sha256": payload_sha})
        if self.token:
            h["x-amz-security-token"] = self.token
        names = sorted(h)
        canonical_q = "&".join(f"{_quote(k)}={_quote(str(v))}" for k, v in sorted(query.items()))
        canonical = "\n".join([method, path, canonical_q, "".join(f"{k}:{h[k]}\n" for k in names), ";".join(names), payload_sha])
        scope = f"{day}/{self.region}/s3/aws4_request"
        to_sign = "\n".join(["AWS4-HMAC-SHA256", amz_date, scope, hashlib.sha256(canonical.encode()).hexdigest()])
        key = ("AWS4" + self.secret_key).encode()
        for part in (day, self.region, "s3", "aws4_request"):
            key = hmac.new(key, part.encode(), hashlib.sha256).digest()
        sig = hmac.new(key, to_sign.encode(), hashlib.sha256).hexdigest()
        h["authorization"] = (f"AWS4-HMAC-SHA256 Credential={self.access_key}/{scope}, SignedHeaders={';'.join(names)}, "
                              f"Signature={sig}")
        return h

    def _request(self, method: str, key: Optional[str], query: Optional[Dict[str, str]] = None, body: Any = None,
                 length: int = 0, payload_sha: str = _EMPTY_SHA, headers: Optional[Dict[str, str]] = None,
                 sink: Any = None, ok: Tuple[int, ...] = (200, 204)) -> Tuple[int, Dict[str, str], bytes]:
        path = f"{self.base}/{_quote(self.bucket)}" + (f"/{_quote(key, safe='-_.~/')}" if key is not None else "")
        query = dict(query or {})
        last: Optional[Exception] = None
        for attempt in range(4):
            hdrs = self._sign(method, path, query, dict(headers or {}), payload_sha)
            if body is not None or method in ("PUT", "POST"):
                hdrs["content-length"] = str(length)
            qs = "&".join(f"{_quote(k)}={_quote(str(v))}" for k, v in sorted(query.items()))
            conn = _connect(self.scheme, self.host, self.timeout)
            src: Any = None
            try:
                src = body() if callable(body) else body
                conn.request(method, path + ("?" + qs if qs else ""), body=src, headers=hdrs)
                resp = conn.getresponse()
                if resp.status >= 500 and attempt < 3:
                    resp.read()
                    time.sleep(0.2 * (2 ** attempt))
                    continue
                out = b""
                if sink is not None and resp.status in ok:
                    while True:
                        chunk = resp.read(1 << 20)
                        if not chunk:
                            break
                        sink.write(chunk)
                else:
                    out = resp.read()
                return resp.status, {k.lower(): v for k, v in resp.getheaders()}, out
            except (OSError, http.client.HTTPException) as e:
                last = e
                time.sleep(0.2 * (2 ** attempt))
            finally:
                if hasattr(src, "close") and callable(body):
                    src.close()
                conn.close()
        raise ObjectStoreError(0, str(last), f"s3 {method} {key}")

    def _check(self, status: int, body: bytes, what: str, ok: Tuple[int, ...] = (200, 204)) -> None:
        if status not in ok:
            raise ObjectStoreError(status, body.decode("utf-8", "replace"), what)

    # -- calls ---------------------------------------------------------------------------------------------------
    def put_bytes(self, key: str, data: bytes) -> None:
        st, _, out = self._request("PUT", key, body=data, length=len(data), payload_sha=hashlib.sha256(data).hexdigest())
        self._check(st, out, f"s3 put {key}")

    def put_file(self, key: str, path: str) -> None:
        size = os.path.getsize(path)
        if size <= MULTIPART_THRESHOLD:
            st, _, out = self._request("PUT", key, body=lambda: open(path, "rb"), length=size, payload_sha=_file_sha256(path))
            self._check(st, out, f"s3 put {key}")
            return
        st, _, out = self._request("POST", key, query={"uploads": ""})
        self._check(st, out, f"s3 start multipart {key}")
        upload_id = _xml_find(out, "UploadId")
        etags: List[Tuple[int, str]] = []
        try:
            for n, off in enumerate(range(0, size, PART_SIZE), 1):
                ln = min(PART_SIZE, size - off)
                st, hd, out = self._request("PUT", key, query={"partNumber": str(n), "uploadId": upload_id},
                                            body=lambda off=off, ln=ln: _FileSlice(path, off, ln), length=ln,
                                            payload_sha=_file_sha256(path, off, ln))
                self._check(st, out, f"s3 upload part {n} of {key}")
                etags.append((n, hd.get("etag", "")))
            doc = "<CompleteMultipartUpload>" + "".join(
                f"<Part><PartNumber>{n}</PartNumber><ETag>{e}</ETag></Part>" for n, e in etags) + "</CompleteMultipartUpload>"
            data = doc.encode()
            st, _, out = self._request("POST", key, query={"uploadId": upload_id}, body=data, length=len(data),
                                       payload_sha=hashlib.sha256(data).hexdigest())
            self._check(st, out, f"s3 complete multipart {key}")
            if b"<Error>" in out:
                raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"s3 complete multipart {key}")
        except Exception:
            self._request("DELETE", key, query={"uploadId": upload_id})
            raise

    def get_file(self, key: str, path: str) -> None:
        tmp = f"{path}.part{os.getpid()}"
        with open(tmp, "wb") as f:
            st, _, out = self._request("GET", key, sink=f)
        if st != 200:
            os.unlink(tmp)
            raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"s3 get {key}")
        os.replace(tmp, path)

    def get_bytes(self, key: str) -> bytes:
        st, _, out = self._request("GET", key)
        self._check(st, out, f"s3 get {key}", ok=(200,))
        return out

    def exists(self, key: str) -> bool:
        st, _, out = self._request("HEAD", key)
        if st == 404:
            return False
        self._check(st, out, f"s3 head {key}", ok=(200,))
        return True

    def delete(self, key: str) -> None:
        st, _, out = self._request("DELETE", key)
        self._check(st, out, f"s3 delete {key}", ok=(200, 204, 404))

    def list(self, prefix: str = "") -> Iterator[Tuple[str, int]]:
        token: Optional[str] = None
        while True:
            q = {"list-type": "2", "prefix": prefix}
            if token:
                q["continuation-token"] = token
            st, _, out = self._request("GET", None, query=q)
            self._check(st, out, f"s3 list {prefix}", ok=(200,))
            root = ET.fromstring(out)
            for el in root.iter():
                if el.tag.endswith("Contents"):
                    vals = {c.tag.rsplit("}", 1)[-1]: c.text for c in el}
                    yield vals.get("Key", ""), int(vals.get("Size") or 0)
            token = _xml_find(out, "NextContinuationToken")
            if not token or _xml_find(out, "IsTruncated") != "true":
                return


def _xml_find(doc: bytes, tag: str) -> str:
    try:
        for el in ET.fromstring(doc).iter():
            if el.tag.rsplit("}", 1)[-1] == tag:
                return el.text or ""
    except ET.ParseError:
        pass
    return ""


class GCSClient:
    """Google Cloud Storage through its JSON API. Authorisation, first that applies: ``STORAGE_EMULATOR_HOST`` (no
    auth, the emulator convention), ``GOOGLE_OAUTH_ACCESS_TOKEN``, a service-account key file named by
    ``GOOGLE_APPLICATION_CREDENTIALS`` (RS256 JWT grant), the GCE metadata server."""

    SCOPE = "https://www.googleapis.com/auth/devstorage.read_write"

    def __init__(self, bucket: str, timeout: float = 120.0) -> None:
        self.bucket, self.timeout = bucket, timeout
        emu = os.environ.get("STORAGE_EMULATOR_HOST")
        if emu:
            url = urllib.parse.urlparse(emu if "://" in emu else "http://" + emu)
            self.scheme, self.host, self.auth = url.scheme, url.netloc, False
        else:
            self.scheme, self.host, self.auth = "https", "storage.googleapis.com", True
        self._token: Optional[str] = None
        self._token_expiry = 0.0

    # -- authorisation -------------------------------------------------------------------------------------------
    def _bearer(self) -> Optional[str]:
        if not self.auth:
            return None
        if self._token and time.time() < self._token_expiry - 60:
            return self._token
        env_tok = os.environ.get("GOOGLE_OAUTH_ACCESS_TOKEN")
        if env_tok:
            self._token, self._token_expiry = env_tok, time.time() + 600
            return self._token
        key_file = os.environ.get("GOOGLE_APPLICATION_CREDENTIALS")
        if key_file and os.path.exists(key_file):
            tok = self._service_account_token(json.load(open(key_file)))
        else:
            tok = self._metadata_token()
        self._token, self._token_expiry = tok["access_token"], time.time() + float(tok.get("expires_in", 300))
        return self._token

    def _service_account_token(self, info: Dict[str, Any]) -> Dict[str, Any]:
        from cryptography.hazmat.primitives import hashes, serialization
        from cryptography.hazmat.primitives.asymmetric import padding

        def b64(b: bytes) -> bytes:
            return base64.urlsafe_b64encode(b).rstrip(b"=")

        token_uri = info.get("token_uri", "https://oauth2.googleapis.com/token")
        now = int(time.time())
        claims = {"iss": info["client_email"], "scope": self.SCOPE, "aud": token_uri, "iat": now, "exp": now + 3600}
        signing_input = b64(json.dumps({"alg": "RS256", "typ": "JWT"}).encode()) + b"." + b64(json.dumps(claims).encode())
        key = serialization.load_pem_private_key(info["private_key"].encode(), password=None)
        jwt = signing_input + b"." + b64(key.sign(signing_input, padding.PKCS1v15(), hashes.SHA256()))
        body = urllib.parse.urlencode({"grant_type": "urn:ietf:params:oauth:grant-type:jwt-bearer", "assertion": jwt.decode()}).encode()
        url = urllib.parse.urlparse(token_uri)
        conn = _connect(url.scheme, url.netloc, self.timeout)
        try:
            conn.request("POST", url.path, body=body, headers={"Content-Type": "application/x-www-form-urlencoded"})
            resp = conn.getresponse()
            out = resp.read()
            if resp.status != 200:
                raise ObjectStoreError(resp.status, out.decode("utf-8", "replace"), "gcs token grant")
            return json.loads(out)
        finally:
            conn.close()

    def _metadata_token(self) -> Dict[str, Any]:
        conn = http.client.HTTPConnection(os.environ.get("GCE_METADATA_HOST", "metadata.google.internal"), timeout=5)
        try:
            conn.request("GET", "/computeMetadata/v1/instance/service-accounts/default/token", headers={"Metadata-Flavor": "Google"})
            resp = conn.getresponse()
            out = resp.read()
            if resp.status != 200:
                raise ObjectStoreError(resp.status, out.decode("utf-8", "replace"), "gcs metadata token")
            return json.loads(out)
        except OSError as e:
            raise RuntimeError("GCS needs credentials: GOOGLE_APPLICATION_CREDENTIALS, GOOGLE_OAUTH_ACCESS_TOKEN, or a GCE "
                               f"service account ({e})") from e
        finally:
            conn.close()

    # -- transport -----------------------------------------------------------------------------------------------
    def _request(self, method: str, path: str, query: Optional[Dict[str, str]] = None, body: Any = None, length: int = 0,
                 sink: Any = None, content_type: str = "application/octet-stream") -> Tuple[int, bytes]:
        qs = urllib.parse.urlencode(query or {})
        last: Optional[Exception] = None
        for attempt in range(4):
            hdrs: Dict[str, str] = {}
            tok = self._bearer()
            if tok:
                hdrs["Authorization"] = "Bearer " + tok
            if body is not None:
                hdrs["Content-Length"], hdrs["Content-Type"] = str(length), content_type
            conn = _connect(self.scheme, self.host, self.timeout)
            src: Any = None
            try:
                src = body() if callable(body) else body
                conn.request(method, path + ("?" + qs if qs else ""), body=src, headers=hdrs)
                resp = conn.getresponse()
                if (resp.status >= 500 or resp.status == 429) and attempt < 3:
                    resp.read()
                    time.sleep(0.2 * (2 ** attempt))
                    continue
                if resp.status == 401 and attempt < 3:
                    resp.read()
                    self._token = None
                    continue
                out = b""
                if sink is not None and resp.status == 200:
                    while True:
                        chunk = resp.read(1 << 20)
                        if not chunk:
                            break
                        sink.write(chunk)
                else:
                    out = resp.read()
                return resp.status, out
            except (OSError, http.client.HTTPException) as e:
                last = e
                time.sleep(0.2 * (2 ** attempt))
            finally:
                if callable(body) and hasattr(src, "close"):
                    src.close()
                conn.close()
        raise ObjectStoreError(0, str(last), f"gcs {method} {path}")

    def _obj(self, key: str) -> str:
        return f"/storage/v1/b/{_quote(self.bucket)}/o/{_quote(key, safe='')}"

    # -- calls ---------------------------------------------------------------------------------------------------
    def put_bytes(self, key: str, data: bytes) -> None:
        st, out = self._request("POST", f"/upload/storage/v1/b/{_quote(self.bucket)}/o", {"uploadType": "media", "name": key},
                                body=data, length=len(data))
        if st != 200:
            raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"gcs put {key}")

    def put_file(self, key: str, path: str) -> None:
        st, out = self._request("POST", f"/upload/storage/v1/b/{_quote(self.bucket)}/o", {"uploadType": "media", "name": key},
                                body=lambda: open(path, "rb"), length=os.path.getsize(path))
        if st != 200:
            raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"gcs put {key}")

    def get_file(self, key: str, path: str) -> None:
        tmp = f"{path}.part{os.getpid()}"
        with open(tmp, "wb") as f:
            st, out = self._request("GET", self._obj(key), {"alt": "media"}, sink=f)
        if st != 200:
            os.unlink(tmp)
            raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"gcs get {key}")
        os.replace(tmp, path)

    def get_bytes(self, key: str) -> bytes:
        st, out = self._request("GET", self._obj(key), {"alt": "media"})
        if st != 200:
            raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"gcs get {key}")
        return out

    def exists(self, key: str) -> bool:
        st, out = self._request("GET", self._obj(key))
        if st == 404:
            return False
        if st != 200:
            raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"gcs stat {key}")
        return True

    def delete(self, key: str) -> None:
        st, out = self._request("DELETE", self._obj(key))
        if st not in (200, 204, 404):
            raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"gcs delete {key}")

    def list(self, prefix: str = "") -> Iterator[Tuple[str, int]]:
        page: Optional[str] = None
        while True:
            q = {"prefix": prefix}
            if page:
                q["pageToken"] = page
            st, out = self._request("GET", f"/storage/v1/b/{_quote(self.bucket)}/o", q)
            if st != 200:
                raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"gcs list {prefix}")
            doc = json.loads(out or b"{}")
            for it in doc.get("items", []):
                yield it["name"], int(it.get("size", 0))
            page = doc.get("nextPageToken")
            if not page:
                return


class WebHDFSClient:
    """HDFS through the NameNode's WebHDFS REST interface (``<hdfs_url>/webhdfs/v1/<path>?op=…``): CREATE and OPEN are
    two-step — the NameNode answers 307 with the DataNode location, the data goes to / comes from there."""

    def __init__(self, url: str, root: str, user: Optional[str] = None, timeout: float = 120.0) -> None:
        u = urllib.parse.urlparse(url)
        self.scheme, self.host, self.root, self.user, self.timeout = u.scheme or "http", u.netloc, "/" + root.strip("/"), user, timeout

    def _path(self, key: str) -> str:
        return "/webhdfs/v1" + _quote(f"{self.root}/{key}".rstrip("/"), safe="-_.~/")

    def _call(self, method: str, scheme: str, host: str, target: str, body: Any = None, length: int = 0,
              sink: Any = None) -> Tuple[int, Dict[str, str], bytes]:
        conn = _connect(scheme, host, self.timeout)
        src: Any = None
        try:
            hdrs = {"Content-Length": str(length), "Content-Type": "application/octet-stream"} if body is not None or method == "PUT" else {}
            src = body() if callable(body) else body
            conn.request(method, target, body=src, headers=hdrs)
            resp = conn.getresponse()
            out = b""
            if sink is not None and resp.status == 200:
                while True:
                    chunk = resp.read(1 << 20)
                    if not chunk:
                        break
                    sink.write(chunk)
            else:
                out = resp.read()
            return resp.status, {k.lower(): v for k, v in resp.getheaders()}, out
        finally:
            if callable(body) and hasattr(src, "close"):
                src.close()
            conn.close()

    def _op(self, method: str, key: str, op: str, body: Any = None, length: int = 0, sink: Any = None,
            **params: str) -> Tuple[int, bytes]:
        q = {"op": op, **params}
        if self.user:
            q["user.name"] = self.user
        target = self._path(key) + "?" + urllib.parse.urlencode(q)
        two_step = op in ("CREATE", "OPEN")
        st, hd, out = self._call(method, self.scheme, self.host, target, None if two_step else body, 0 if two_step else length,
                                 None if two_step else sink)
        if st in (307, 302) and "location" in hd:
            loc = urllib.parse.urlparse(hd["location"])
            st, hd, out = self._call(method, loc.scheme or self.scheme, loc.netloc, loc.path + ("?" + loc.query if loc.query else ""),
                                     body, length, sink)
        elif two_step and st == 200 and op == "OPEN":
            st, hd, out = self._call(method, self.scheme, self.host, target, None, 0, sink)   # a gateway that serves directly
        return st, out

    def mkdirs(self, key: str) -> None:
        st, out = self._op("PUT", key, "MKDIRS")
        if st != 200:
            raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"hdfs mkdirs {key}")

    def put_bytes(self, key: str, data: bytes) -> None:
        st, out = self._op("PUT", key, "CREATE", body=data, length=len(data), overwrite="true")
        if st not in (200, 201):
            raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"hdfs create {key}")

    def put_file(self, key: str, path: str) -> None:
        st, out = self._op("PUT", key, "CREATE", body=lambda: open(path, "rb"), length=os.path.getsize(path), overwrite="true")
        if st not in (200, 201):
            raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"hdfs create {key}")

    def get_file(self, key: str, path: str) -> None:
        tmp = f"{path}.part{os.getpid()}"
        with open(tmp, "wb") as f:
            st, out = self._op("GET", key, "OPEN", sink=f)
        if st != 200:
            os.unlink(tmp)
            raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"hdfs open {key}")
        os.replace(tmp, path)

    def exists(self, key: str) -> bool:
        st, _ = self._op("GET", key, "GETFILESTATUS")
        return st == 200

    def delete(self, key: str, recursive: bool = True) -> None:
        st, out = self._op("DELETE", key, "DELETE", recursive="true" if recursive else "false")
        if st not in (200, 404):
            raise ObjectStoreError(st, out.decode("utf-8", "replace"), f"hdfs delete {key}")
