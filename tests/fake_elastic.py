"""A stand-in Elasticsearch for the trial-log backend test: `_bulk` (index actions) and `<index>/_search` with a term
filter, a sort field and from / size — the slice of the API csrc/core/elastic_logs.hpp uses."""
import json
import sys
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

DOCS = {}
LOCK = threading.Lock()


class H(BaseHTTPRequestHandler):
    def log_message(self, *a):
        pass

    def _body(self):
        return self.rfile.read(int(self.headers.get("Content-Length", "0")))

    def _send(self, obj, code=200):
        data = json.dumps(obj).encode()
        self.send_response(code)
        self.send_header("Content-Type", "application/json")
        self.send_header("Content-Length", str(len(data)))
        self.end_headers()
        self.wfile.write(data)

    def do_POST(self):
        path = self.path.split("?")[0]
        if path == "/_bulk":
            lines = [l for l in self._body().decode().split("\n") if l.strip()]
            items = []
            with LOCK:
                for action, doc in zip(lines[0::2], lines[1::2]):
                    idx = json.loads(action)["index"]["_index"]
                    DOCS.setdefault(idx, []).append(json.loads(doc))
                    items.append({"index": {"_index": idx, "status": 201}})
            return self._send({"errors": False, "items": items})
        if path.endswith("/_search"):
            idx = path.strip("/").split("/")[0]
            q = json.loads(self._body() or b"{}")
            with LOCK:
                docs = list(DOCS.get(idx, []))
            for k, v in (q.get("query", {}).get("term") or {}).items():
                docs = [d for d in docs if d.get(k) == v]
            for s in reversed(q.get("sort", [])):
                (field, order), = s.items()
                docs.sort(key=lambda d: d.get(field), reverse=(order == "desc"))
            frm, size = int(q.get("from", 0)), int(q.get("size", 10))
            return self._send({"hits": {"total": {"value": len(docs)}, "hits": [{"_index": idx, "_source": d} for d in docs[frm:frm + size]]}})
        self._send({"error": "unsupported"}, 400)

    def do_GET(self):
        if self.path.startswith("/_count"):
            with LOCK:
                return self._send({"count": sum(len(v) for v in DOCS.values()), "indices": {k: len(v) for k, v in DOCS.items()}})
        self._send({"name": "fake-es", "version": {"number": "7.9.0"}})


if __name__ == "__main__":
    srv = ThreadingHTTPServer(("127.0.0.1", int(sys.argv[1]) if len(sys.argv) > 1 else 0), H)
    print(srv.server_address[1], flush=True)
    srv.serve_forever()
