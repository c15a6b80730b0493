"""In-process stand-ins for the three remote stores the storage layer talks to: S3 (checks the SigV4 signature of every
request with its own implementation of the algorithm), the GCS JSON API, and WebHDFS (NameNode redirect + DataNode)."""
import hashlib
import hmac
import json
import threading
import urllib.parse
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Dict, Tuple


def _serve(handler_cls) -> Tuple[ThreadingHTTPServer, str]:
    srv = ThreadingHTTPServer(("127.0.0.1", 0), handler_cls)
    srv.daemon_threads = True
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    return srv, f"http://127.0.0.1:{srv.server_address[1]}"


class _Base(BaseHTTPRequestHandler):
    protocol_version = "HTTP/1.1"

    def log_message(self, *a):
        pass

    def _body(self) -> bytes:
        n = int(self.headers.get("Content-Length") or 0)
        return self.rfile.read(n) if n else b""

    def _send(self, status: int, body: bytes = b"", headers: Dict[str, str] = None) -> None:
        self.send_response(status)
        for k, v in (headers or {}).items():
            self.send_header(k, v)
        self.send_header("Content-Length", str(len(body)))
        self.end_headers()
        if self.command != "HEAD":
            self.wfile.write(body)


# ---- S3 ------------------------------------------------------------------------------------------------------------------
def make_s3(access_key: str = "AKIDEXAMPLE", secret_key: str = "s3cr3t", region: str = "us-east-1", page: int = 2):
    objects: Dict[str, bytes] = {}
    uploads: Dict[str, Dict[int, bytes]] = {}
    stats = {"requests": 0, "bad_signature": 0, "parts": 0}

    def uri_encode(s: str, slash_ok: bool) -> str:
        return urllib.parse.quote(s, safe="-_.~" + ("/" if slash_ok else ""))

    class H(_Base):
        def _verify(self, body: bytes) -> bool:
            auth = self.headers.get("Authorization", "")
            try:
                fields = dict(p.strip().split("=", 1) for p in auth.split(" ", 1)[1].split(","))
                cred, signed, sig = fields["Credential"], fields["SignedHeaders"].split(";"), fields["Signature"]
                key_id, day, reg, svc, _ = cred.split("/")
            except Exception:
                return False
            url = urllib.parse.urlsplit(self.path)
            q = urllib.parse.parse_qsl(url.query, keep_blank_values=True)
            cq = "&".join(f"{uri_encode(k, False)}={uri_encode(v, False)}" for k, v in sorted(q))
            ch = "".join(f"{h}:{' '.join(self.headers.get(h, '').split())}\n" for h in signed)
            payload = self.headers.get("x-amz-content-sha256", "")
            if payload != hashlib.sha256(body).hexdigest():
                return False
            canonical = "\n".join([self.command, url.path, cq, ch, ";".join(signed), payload])
            to_sign = "\n".join(["AWS4-HMAC-SHA256", self.headers.get("x-amz-date", ""), f"{day}/{reg}/{svc}/aws4_request",
                                 hashlib.sha256(canonical.encode()).hexdigest()])
            k = ("AWS4" + secret_key).encode()
            for part in (day, reg, svc, "aws4_request"):
                k = hmac.new(k, part.encode(), hashlib.sha256).digest()
            return key_id == access_key and reg == region and hmac.new(k, to_sign.encode(), hashlib.sha256).hexdigest() == sig

        def _handle(self) -> None:
            body = self._body()
            stats["requests"] += 1
            if not self._verify(body):
                stats["bad_signature"] += 1
                return self._send(403, b"<Error><Code>SignatureDoesNotMatch</Code></Error>")
            url = urllib.parse.urlsplit(self.path)
            q = dict(urllib.parse.parse_qsl(url.query, keep_blank_values=True))
            parts = urllib.parse.unquote(url.path).lstrip("/").split("/", 1)
            key = parts[1] if len(parts) > 1 else None
            m = self.command
            if key is None and m == "GET":            # ListObjectsV2
                keys = sorted(k for k in objects if k.startswith(q.get("prefix", "")))
                start = int(q.get("continuation-token", 0))
                chunk = keys[start:start + page]
                more = start + page < len(keys)
                xml = "<ListBucketResult>" + "".join(f"<Contents><Key>{k}</Key><Size>{len(objects[k])}</Size></Contents>" for k in chunk)
                xml += f"<IsTruncated>{'true' if more else 'false'}</IsTruncated>"
                xml += f"<NextContinuationToken>{start + page}</NextContinuationToken>" if more else ""
                return self._send(200, (xml + "</ListBucketResult>").encode())
            if m == "POST" and "uploads" in q:
                uid = f"up{len(uploads)}"
                uploads[uid] = {}
                return self._send(200, f"<InitiateMultipartUploadResult><UploadId>{uid}</UploadId></InitiateMultipartUploadResult>".encode())
            if m == "PUT" and "uploadId" in q:
                uploads[q["uploadId"]][int(q["partNumber"])] = body
                stats["parts"] += 1
                return self._send(200, b"", {"ETag": f'"{hashlib.md5(body).hexdigest()}"'})
            if m == "POST" and "uploadId" in q:
                ps = uploads.pop(q["uploadId"])
                objects[key] = b"".join(ps[i] for i in sorted(ps))
                return self._send(200, b"<CompleteMultipartUploadResult/>")
            if m == "DELETE" and "uploadId" in q:
                uploads.pop(q["uploadId"], None)
                return self._send(204)
            if m == "PUT":
                objects[key] = body
                return self._send(200)
            if m in ("GET", "HEAD"):
                if key not in objects:
                    return self._send(404, b"<Error><Code>NoSuchKey</Code></Error>")
                return self._send(200, objects[key])
            if m == "DELETE":
                objects.pop(key, None)
                return self._send(204)
            self._send(400)

        do_GET = do_PUT = do_POST = do_DELETE = do_HEAD = _handle

    srv, url = _serve(H)
    return srv, url, objects, stats


# ---- GCS JSON API -----------------------------------------------------------------------------------------------------
def make_gcs(page: int = 2):
    objects: Dict[str, bytes] = {}

    class H(_Base):
        def _handle(self) -> None:
            body = self._body()
            url = urllib.parse.urlsplit(self.path)
            q = dict(urllib.parse.parse_qsl(url.query, keep_blank_values=True))
            segs = url.path.split("/")
            m = self.command
            if url.path.startswith("/upload/storage/v1/b/") and m == "POST" and q.get("uploadType") == "media":
                objects[q["name"]] = body
                return self._send(200, json.dumps({"name": q["name"], "size": str(len(body))}).encode())
            if url.path.startswith("/storage/v1/b/") and len(segs) == 6 and segs[5] == "o" and m == "GET":
                keys = sorted(k for k in objects if k.startswith(q.get("prefix", "")))
                start = int(q.get("pageToken", 0))
                doc = {"items": [{"name": k, "size": str(len(objects[k]))} for k in keys[start:start + page]]}
                if start + page < len(keys):
                    doc["nextPageToken"] = str(start + page)
                return self._send(200, json.dumps(doc).encode())
            if url.path.startswith("/storage/v1/b/") and len(segs) == 7 and segs[5] == "o":
                key = urllib.parse.unquote(segs[6])
                if m == "DELETE":
                    return self._send(204 if objects.pop(key, None) is not None else 404)
                if key not in objects:
                    return self._send(404, b'{"error": {"code": 404}}')
                if q.get("alt") == "media":
                    return self._send(200, objects[key])
                return self._send(200, json.dumps({"name": key, "size": str(len(objects[key]))}).encode())
            self._send(400)

        do_GET = do_POST = do_DELETE = _handle

    srv, url = _serve(H)
    return srv, url, objects


# ---- WebHDFS ------------------------------------------------------------------------------------------------------------
def make_webhdfs():
    files: Dict[str, bytes] = {}
    dirs = set()
    stats = {"redirects": 0}

    class H(_Base):
        def _handle(self) -> None:
            body = self._body()
            url = urllib.parse.urlsplit(self.path)
            q = dict(urllib.parse.parse_qsl(url.query, keep_blank_values=True))
            assert url.path.startswith("/webhdfs/v1/")
            path = urllib.parse.unquote(url.path[len("/webhdfs/v1"):])
            op, m = q.get("op"), self.command
            if op in ("CREATE", "OPEN") and "datanode" not in q:
                stats["redirects"] += 1
                host = self.headers.get("Host")
                return self._send(307, b"", {"Location": f"http://{host}{url.path}?{url.query}&datanode=true"})
            if op == "CREATE" and m == "PUT":
                files[path] = body
                return self._send(201)
            if op == "OPEN" and m == "GET":
                return self._send(200, files[path]) if path in files else self._send(404, b"{}")
            if op == "MKDIRS" and m == "PUT":
                dirs.add(path)
                return self._send(200, b'{"boolean": true}')
            if op == "GETFILESTATUS":
                return self._send(200 if path in files or path in dirs else 404, b"{}")
            if op == "DELETE" and m == "DELETE":
                gone = [p for p in files if p == path or p.startswith(path + "/")] if q.get("recursive") == "true" else [p for p in files if p == path]
                for p in gone:
                    del files[p]
                for d in [d for d in dirs if d == path or d.startswith(path + "/")]:
                    dirs.discard(d)
                return self._send(200, b'{"boolean": true}')
            self._send(400)

        do_GET = do_PUT = do_DELETE = _handle

    srv, url = _serve(H)
    return srv, url, files, stats
