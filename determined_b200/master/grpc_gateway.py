"""gRPC front of the master: serves ``determined.api.v1.Determined`` and translates every RPC to the REST route that
``api_spec.RPCS`` binds it to.

The reference's master speaks gRPC natively and adds REST through grpc-gateway (``master/internal/grpc/api.go:1-101``,
``proto/src/determined/api/v1/api.proto``). This build's master is a C++ REST server, so the direction is reversed: a
small sidecar turns gRPC into REST. Message classes come from descriptors built out of the API table at start-up — the
same table the published ``.proto`` is generated from — so no generated ``_pb2`` modules and no ``protoc`` are needed.

    python -m determined_b200.master.grpc_gateway --master 127.0.0.1:8080 --port 8090

Authentication: the bearer token of ``Login`` travels in the ``authorization`` metadata key (``Bearer <token>``) or in
``x-user-token`` (the reference's key) and is forwarded to the master. REST errors map to gRPC status codes
(404 NOT_FOUND, 400 INVALID_ARGUMENT, 401 UNAUTHENTICATED, 403 PERMISSION_DENIED, 409 FAILED_PRECONDITION).
"""
from __future__ import annotations

import argparse
import json
import time
import urllib.error
import urllib.parse
import urllib.request
from concurrent import futures
from typing import Any, Dict, Iterator, List, Optional, Tuple

from . import api_spec as spec

_SCALARS = {"string": 9, "int32": 5, "int64": 3, "double": 1, "bool": 8}   # FieldDescriptorProto.Type values
_WELL_KNOWN = {"Struct": ".google.protobuf.Struct", "Value": ".google.protobuf.Value"}


def _add_fields(msg: Any, fields: List[Tuple[str, ...]], with_raw: bool) -> None:
    for number, f in enumerate(fields, 1):
        fd = msg.field.add()
        fd.name, fd.number = f[0], number
        fd.label = 3 if len(f) > 2 and f[2] == "repeated" else 1
        if f[1] in _SCALARS:
            fd.type = _SCALARS[f[1]]
        else:
            fd.type = 11
            fd.type_name = _WELL_KNOWN.get(f[1], f".{spec.PACKAGE}.{f[1]}")
    if with_raw:
        fd = msg.field.add()
        fd.name, fd.number, fd.label, fd.type, fd.type_name = "raw", 100, 1, 11, _WELL_KNOWN["Struct"]


def file_descriptor_proto() -> Any:
    from google.protobuf import descriptor_pb2

    fdp = descriptor_pb2.FileDescriptorProto(name="determined/api/v1/api.proto", package=spec.PACKAGE, syntax="proto3")
    fdp.dependency.append("google/protobuf/struct.proto")
    for name, fields in spec.RESOURCES.items():
        m = fdp.message_type.add()
        m.name = name
        _add_fields(m, fields, with_raw=True)
    svc = fdp.service.add()
    svc.name = spec.SERVICE
    for rpc in spec.RPCS:
        req = fdp.message_type.add()
        req.name = rpc["name"] + "Request"
        _add_fields(req, rpc["request"], with_raw=False)
        resp = fdp.message_type.add()
        resp.name = rpc["name"] + "Response"
        if rpc["response"] is not None:
            _add_fields(resp, [rpc["response"]], with_raw=False)
        meth = svc.method.add()
        meth.name = rpc["name"]
        meth.input_type = f".{spec.PACKAGE}.{req.name}"
        meth.output_type = f".{spec.PACKAGE}.{resp.name}"
        meth.server_streaming = bool(rpc["stream"])
    return fdp


_CLASSES: Dict[str, Any] = {}


def message_classes() -> Dict[str, Any]:
    """``{message name: class}`` for every message of the API (built once per process)."""
    if _CLASSES:
        return _CLASSES
    from google.protobuf import descriptor_pool, message_factory, struct_pb2

    pool = descriptor_pool.DescriptorPool()
    pool.AddSerializedFile(struct_pb2.DESCRIPTOR.serialized_pb)
    fdp = file_descriptor_proto()
    pool.Add(fdp)
    for m in fdp.message_type:
        _CLASSES[m.name] = message_factory.GetMessageClass(pool.FindMessageTypeByName(f"{spec.PACKAGE}.{m.name}"))
    return _CLASSES


def _fill(msg: Any, fields: List[Tuple[str, ...]], obj: Dict[str, Any], with_raw: bool) -> None:
    """Typed fields are filled one by one and a value that does not fit its declared type is skipped (it is still in ``raw``)."""
    from google.protobuf import json_format

    for f in fields:
        if f[0] not in obj or obj[f[0]] is None:
            continue
        v = obj[f[0]]
        repeated = len(f) > 2 and f[2] == "repeated"
        try:
            if f[1] in spec.RESOURCES:
                cls = message_classes()[f[1]]
                items = v if repeated else [v]
                for item in items:
                    sub = cls()
                    if isinstance(item, dict):
                        _fill(sub, spec.RESOURCES[f[1]], item, with_raw=True)
                    if repeated:
                        getattr(msg, f[0]).append(sub)
                    else:
                        getattr(msg, f[0]).CopyFrom(sub)
            else:
                if f[1] in ("int32", "int64") and isinstance(v, float) and v == int(v):
                    v = int(v)
                json_format.ParseDict({f[0]: v}, msg, ignore_unknown_fields=True)
        except Exception:
            msg.ClearField(f[0])
    if with_raw:
        try:
            msg.raw.update(obj)
        except Exception:
            pass


def _intify(x: Any) -> Any:
    if isinstance(x, float) and x.is_integer() and abs(x) < 2 ** 53:
        return int(x)
    if isinstance(x, dict):
        return {k: _intify(v) for k, v in x.items()}
    if isinstance(x, list):
        return [_intify(v) for v in x]
    return x


class RestBackend:
    def __init__(self, master: str) -> None:
        from ..utils import tls

        if "://" not in master:
            master = "http://" + master
        self.base = master.rstrip("/")
        self._opener = tls.build_opener()

    def call(self, method: str, path: str, body: Any, token: Optional[str], timeout: float = 70.0) -> Any:
        headers = {"Content-Type": "application/json"}
        if token:
            headers["Authorization"] = token if token.lower().startswith("bearer ") else f"Bearer {token}"
        data = None if body is None else json.dumps(body).encode()
        req = urllib.request.Request(self.base + path, data=data, method=method, headers=headers)
        with self._opener.open(req, timeout=timeout) as resp:
            text = resp.read().decode()
            return json.loads(text) if text else None


def _status_for(code: int) -> Any:
    import grpc

    return {400: grpc.StatusCode.INVALID_ARGUMENT, 401: grpc.StatusCode.UNAUTHENTICATED, 403: grpc.StatusCode.PERMISSION_DENIED,
            404: grpc.StatusCode.NOT_FOUND, 405: grpc.StatusCode.UNIMPLEMENTED, 409: grpc.StatusCode.FAILED_PRECONDITION}.get(
        code, grpc.StatusCode.INTERNAL if code >= 500 else grpc.StatusCode.UNKNOWN)


class Gateway:
    def __init__(self, master: str) -> None:
        self.backend = RestBackend(master)
        self.classes = message_classes()

    # ---- request side ----------------------------------------------------------------------------------------------------
    def _rest_request(self, rpc: Dict[str, Any], req: Any) -> Tuple[str, Any]:
        from google.protobuf import json_format

        # google.protobuf.Struct carries every number as a double: give integral values back their JSON integer form, or
        # `global_batch_size: 8.0` and file modes fail the master's type checks
        d = _intify(json_format.MessageToDict(req, preserving_proto_field_name=True))
        path = rpc["path"]
        for name in spec.path_fields(rpc):
            value = d.pop(name, None)
            if value in (None, ""):
                value = getattr(req, name)
            path = path.replace("{" + name + "}", urllib.parse.quote(str(value), safe=""))
        if rpc["method"] in ("GET", "DELETE"):
            q = {k: (str(v).lower() if isinstance(v, bool) else v) for k, v in d.items() if not isinstance(v, (dict, list))}
            return path + ("?" + urllib.parse.urlencode(q) if q else ""), None
        if rpc["body"] != "fields":
            return path, d.get(rpc["body"], {})
        return path, d

    def _token(self, context: Any) -> Optional[str]:
        md = dict(context.invocation_metadata() or ())
        return md.get("authorization") or md.get("x-user-token")

    def _response(self, rpc: Dict[str, Any], payload: Any) -> Any:
        resp = self.classes[rpc["name"] + "Response"]()
        f = rpc["response"]
        if f is not None and payload is not None:
            _fill(resp, [f], {f[0]: payload}, with_raw=False)
        return resp

    def _abort(self, context: Any, e: Exception) -> None:
        import grpc

        if isinstance(e, urllib.error.HTTPError):
            text = e.read().decode(errors="replace")
            try:
                text = json.loads(text).get("error", text)
            except Exception:
                pass
            context.abort(_status_for(e.code), text)
        context.abort(grpc.StatusCode.UNAVAILABLE, f"master unreachable: {e!r}")

    def unary(self, rpc: Dict[str, Any]):
        def handler(req: Any, context: Any) -> Any:
            try:
                path, body = self._rest_request(rpc, req)
                return self._response(rpc, self.backend.call(rpc["method"], path, body, self._token(context)))
            except Exception as e:  # noqa: BLE001 - every failure becomes a gRPC status
                self._abort(context, e)

        return handler

    def streaming(self, rpc: Dict[str, Any]):
        """List-returning routes as server streams: one response message per batch. ``TrialLogs`` with ``follow`` keeps polling
        from the last seen entry until the trial reaches a terminal state (the REST route long-polls for new lines)."""

        def handler(req: Any, context: Any) -> Iterator[Any]:
            token = self._token(context)
            try:
                if rpc["name"] != "TrialLogs":
                    path, _ = self._rest_request(rpc, req)
                    items = self.backend.call(rpc["method"], path, None, token) or []
                    if items:
                        yield self._response(rpc, items)
                    return
                # trial logs: the REST route addresses lines by position; filters are applied here so that positions stay
                # meaningful while following
                offset, left = int(req.offset), int(req.limit) if req.limit > 0 else None
                while context.is_active():
                    lines = self.backend.call("GET", f"/api/v1/trials/{req.id}/logs?offset={offset}&limit=5000", None, token) or []
                    offset += len(lines)
                    keep = [l for l in lines if (not req.search or req.search in str(l.get("message", "")))
                            and (not req.level or str(l.get("level", "")) == req.level)]
                    if left is not None:
                        keep, left = keep[:left], left - len(keep[:left])
                    if keep:
                        yield self._response(rpc, keep)
                    if not req.follow or (left is not None and left <= 0):
                        return
                    if not lines:
                        trial = self.backend.call("GET", f"/api/v1/trials/{req.id}", None, token)
                        if trial.get("state") in ("COMPLETED", "CANCELED", "ERROR"):
                            return
                        time.sleep(0.5)
            except Exception as e:  # noqa: BLE001
                self._abort(context, e)

        return handler

    def generic_handler(self) -> Any:
        import grpc

        handlers = {}
        for rpc in spec.RPCS:
            req_cls, resp_cls = self.classes[rpc["name"] + "Request"], self.classes[rpc["name"] + "Response"]
            make = grpc.unary_stream_rpc_method_handler if rpc["stream"] else grpc.unary_unary_rpc_method_handler
            handlers[rpc["name"]] = make(self.streaming(rpc) if rpc["stream"] else self.unary(rpc),
                                         request_deserializer=req_cls.FromString, response_serializer=resp_cls.SerializeToString)
        return grpc.method_handlers_generic_handler(f"{spec.PACKAGE}.{spec.SERVICE}", handlers)


def serve(master: str, port: int = 0, host: str = "127.0.0.1", max_workers: int = 16) -> Tuple[Any, int]:
    """Start the gateway; returns ``(server, bound port)``."""
    import grpc

    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers))
    server.add_generic_rpc_handlers((Gateway(master).generic_handler(),))
    bound = server.add_insecure_port(f"{host}:{port}")
    server.start()
    return server, bound


class Client:
    """Minimal typed client over the same descriptors: ``Client("127.0.0.1:8090").GetMaster()`` → response message.
    Keyword arguments become request fields; dicts / lists are accepted for ``Struct`` fields."""

    def __init__(self, target: str, token: Optional[str] = None) -> None:
        import grpc

        self.channel = grpc.insecure_channel(target)
        self.token = token
        self.classes = message_classes()
        self._rpcs = {r["name"]: r for r in spec.RPCS}

    def __getattr__(self, name: str) -> Any:
        if name.startswith("_") or name not in self._rpcs:
            raise AttributeError(name)
        rpc = self._rpcs[name]
        req_cls, resp_cls = self.classes[name + "Request"], self.classes[name + "Response"]
        path = f"/{spec.PACKAGE}.{spec.SERVICE}/{name}"
        make = self.channel.unary_stream if rpc["stream"] else self.channel.unary_unary
        stub = make(path, request_serializer=req_cls.SerializeToString, response_deserializer=resp_cls.FromString)

        def call(**kwargs: Any) -> Any:
            from google.protobuf import json_format

            req = json_format.ParseDict(kwargs, req_cls())
            md = (("authorization", f"Bearer {self.token}"),) if self.token else ()
            return stub(req, metadata=md, timeout=kwargs.pop("_timeout", 90))

        return call


def main() -> None:
    ap = argparse.ArgumentParser(description="gRPC gateway of det-master")
    ap.add_argument("--master", default="127.0.0.1:8080")
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=8090)
    ap.add_argument("--print-port-file", default="")
    args = ap.parse_args()
    server, port = serve(args.master, args.port, args.host)
    if args.print_port_file:
        with open(args.print_port_file, "w") as f:
            f.write(str(port))
    print(f"grpc gateway: {spec.PACKAGE}.{spec.SERVICE} on {args.host}:{port} -> {args.master}", flush=True)
    server.wait_for_termination()


if __name__ == "__main__":
    main()
