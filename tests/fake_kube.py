"""A stand-in for kube-apiserver + kubelet, enough for the master's Kubernetes back-end: nodes with allocatable GPUs,
pod create / list (label selector) / delete, pod logs. "Scheduling" a pod means running its container command as a local
process with the pod's env; the phase follows the process (Pending → Running → Succeeded / Failed)."""
import json
import os
import subprocess
import tempfile
import threading
import time
import urllib.parse
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer


class FakeKube:
    def __init__(self, nodes=None, start_delay=0.3):
        self.nodes = nodes or {"node-a": 2, "node-b": 2}
        self.pods = {}          # name -> dict(manifest, proc, log_path, created, deleted)
        self.created = []       # manifests in creation order (for assertions)
        self.deleted = []
        self.fail_next_creates = 0
        self.start_delay = start_delay
        self.lock = threading.Lock()
        self.tmp = tempfile.mkdtemp(prefix="fakekube")
        outer = self

        class H(BaseHTTPRequestHandler):
            def log_message(self, *a):
                pass

            def _send(self, code, body, ctype="application/json"):
                data = body if isinstance(body, bytes) else json.dumps(body).encode()
                self.send_response(code)
                self.send_header("Content-Type", ctype)
                self.send_header("Content-Length", str(len(data)))
                self.end_headers()
                self.wfile.write(data)

            def do_GET(self):
                u = urllib.parse.urlparse(self.path)
                parts = [p for p in u.path.split("/") if p]
                if u.path == "/api/v1/nodes":
                    items = [{"metadata": {"name": n}, "spec": {}, "status": {"allocatable": {"nvidia.com/gpu": str(g), "cpu": "64"}}}
                             for n, g in outer.nodes.items()]
                    return self._send(200, {"kind": "NodeList", "items": items})
                if len(parts) == 5 and parts[:3] == ["api", "v1", "namespaces"] and parts[4] == "pods":
                    sel = urllib.parse.parse_qs(u.query).get("labelSelector", [""])[0]
                    return self._send(200, {"kind": "PodList", "items": outer.list_pods(parts[3], sel)})
                if len(parts) == 7 and parts[4] == "pods" and parts[6] == "log":
                    with outer.lock:
                        p = outer.pods.get(parts[5])
                    if p is None:
                        return self._send(404, {"message": "not found"})
                    try:
                        data = open(p["log_path"], "rb").read()
                    except FileNotFoundError:
                        data = b""
                    return self._send(200, data, "text/plain")
                return self._send(404, {"message": "no route " + u.path})

            def do_POST(self):
                u = urllib.parse.urlparse(self.path)
                parts = [p for p in u.path.split("/") if p]
                body = self.rfile.read(int(self.headers.get("Content-Length", 0) or 0))
                if len(parts) == 5 and parts[4] == "pods":
                    with outer.lock:
                        if outer.fail_next_creates > 0:
                            outer.fail_next_creates -= 1
                            return self._send(503, {"message": "try again"})
                    m = json.loads(body)
                    code, out = outer.create_pod(parts[3], m)
                    return self._send(code, out)
                return self._send(404, {"message": "no route"})

            def do_DELETE(self):
                u = urllib.parse.urlparse(self.path)
                parts = [p for p in u.path.split("/") if p]
                if len(parts) == 6 and parts[4] == "pods":
                    return self._send(*outer.delete_pod(parts[5]))
                return self._send(404, {"message": "no route"})

        self.httpd = ThreadingHTTPServer(("127.0.0.1", 0), H)
        self.port = self.httpd.server_address[1]
        self.thread = threading.Thread(target=self.httpd.serve_forever, daemon=True)
        self.thread.start()

    # ---- pod lifecycle -------------------------------------------------------------------------------------------------
    def create_pod(self, ns, manifest):
        name = manifest["metadata"]["name"]
        with self.lock:
            if name in self.pods:
                return 409, {"message": "already exists"}
            c = manifest["spec"]["containers"][0]
            env = dict(os.environ)
            for e in c.get("env", []):
                env[e["name"]] = e["value"]
            log_path = os.path.join(self.tmp, name + ".log")
            rec = {"manifest": manifest, "ns": ns, "proc": None, "log_path": log_path, "created": time.time(), "cmd": c["command"],
                   "env": env, "cwd": c.get("workingDir") or None}
            self.pods[name] = rec
            self.created.append(manifest)
        return 201, manifest

    def _maybe_start(self, rec):
        if rec["proc"] is None and time.time() - rec["created"] >= self.start_delay:
            rec["proc"] = subprocess.Popen(rec["cmd"], env=rec["env"], cwd=rec["cwd"], stdout=open(rec["log_path"], "wb"),
                                           stderr=subprocess.STDOUT, start_new_session=True)

    def list_pods(self, ns, selector):
        want = dict(kv.split("=", 1) for kv in selector.split(",") if "=" in kv)
        out = []
        with self.lock:
            for name, rec in self.pods.items():
                labels = rec["manifest"]["metadata"].get("labels", {})
                if rec["ns"] != ns or any(labels.get(k) != v for k, v in want.items()):
                    continue
                self._maybe_start(rec)
                status = {"phase": "Pending", "conditions": [{"type": "PodScheduled", "status": "True"}]}
                proc = rec["proc"]
                if proc is not None:
                    rc = proc.poll()
                    if rc is None:
                        status = {"phase": "Running", "containerStatuses": [{"state": {"running": {}}}]}
                    else:
                        status = {"phase": "Succeeded" if rc == 0 else "Failed",
                                  "containerStatuses": [{"state": {"terminated": {"exitCode": rc if rc >= 0 else 128 - rc,
                                                                                 "reason": "Completed" if rc == 0 else "Error"}}}]}
                m = dict(rec["manifest"])
                m["status"] = status
                out.append(m)
        return out

    def delete_pod(self, name):
        with self.lock:
            rec = self.pods.pop(name, None)
            if rec is None:
                return 404, {"message": "not found"}
            self.deleted.append(name)
        if rec["proc"] is not None and rec["proc"].poll() is None:
            try:
                os.killpg(rec["proc"].pid, 9)
            except ProcessLookupError:
                pass
        return 200, {"status": "Success"}

    def close(self):
        with self.lock:
            recs = list(self.pods.values())
            self.pods.clear()
        for rec in recs:
            if rec["proc"] is not None and rec["proc"].poll() is None:
                try:
                    os.killpg(rec["proc"].pid, 9)
                except ProcessLookupError:
                    pass
        self.httpd.shutdown()
