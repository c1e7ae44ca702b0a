"""Small REST façade over the in-process API server so `kubedlctl` (and curl) can talk to a
running manager the way kubectl talks to kube-apiserver.  Paths follow the Kubernetes layout:

  GET/POST   /apis/<group>/<version>/namespaces/<ns>/<plural>          list / create
  GET/PUT/DELETE  .../<plural>/<name>[/status]                          get / replace / delete
  GET/POST   /api/v1/namespaces/<ns>/{pods,services,events}            core kinds
  POST       /apply            multi-document YAML or JSON list         create-or-replace
  GET        /logs/<ns>/<pod>                                           pod log from the node agent
  GET        /healthz, /apis                                            liveness, served kinds
"""
from __future__ import annotations

import json
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Callable, Dict, Optional, Tuple

import yaml

from ..api.kinds import SCHEME
from .store import ApiError, APIServer

_CORE = {"pods": "Pod", "services": "Service", "events": "Event", "configmaps": "ConfigMap"}


def _plural_map() -> Dict[Tuple[str, str, str], str]:
    out = {}
    for info in SCHEME.workload_kinds():
        out[(info.group, info.version, info.plural)] = info.kind
    out[("scheduling.incubator.k8s.io", "v1alpha1", "podgroups")] = "PodGroup"
    return out


def make_handler(api: APIServer, log_reader: Optional[Callable[[str, str], str]] = None):
    plural = _plural_map()

    class Handler(BaseHTTPRequestHandler):
        def log_message(self, *a):  # quiet
            pass

        def _send(self, code: int, body) -> None:
            data = body if isinstance(body, bytes) else json.dumps(body).encode()
            self.send_response(code)
            self.send_header("Content-Type", "application/json" if not isinstance(body, bytes) else "text/plain")
            self.send_header("Content-Length", str(len(data)))
            self.end_headers()
            self.wfile.write(data)

        def _body(self):
            n = int(self.headers.get("Content-Length", "0") or 0)
            return self.rfile.read(n) if n else b""

        def _route(self):
            parts = [p for p in self.path.split("?")[0].split("/") if p]
            if not parts:
                return None
            if parts[0] == "api" and len(parts) >= 5 and parts[1] == "v1" and parts[2] == "namespaces":
                kind = _CORE.get(parts[4])
                return (kind, parts[3], parts[5] if len(parts) > 5 else None, len(parts) > 6 and parts[6] == "status")
            if parts[0] == "apis" and len(parts) >= 6 and parts[3] == "namespaces":
                kind = plural.get((parts[1], parts[2], parts[5]))
                return (kind, parts[4], parts[6] if len(parts) > 6 else None, len(parts) > 7 and parts[7] == "status")
            return None

        def do_GET(self):
            try:
                if self.path == "/healthz":
                    return self._send(200, {"status": "ok"})
                if self.path == "/apis":
                    return self._send(200, {"kinds": [k for k in SCHEME.kinds() if api.crd_installed(k)]})
                if self.path.startswith("/logs/"):
                    _, _, ns, pod = self.path.split("/", 3)
                    text = log_reader(ns, pod) if log_reader else ""
                    return self._send(200, text.encode())
                r = self._route()
                if not r or not r[0]:
                    return self._send(404, {"reason": "NotFound", "message": self.path})
                kind, ns, name, _ = r
                if name:
                    return self._send(200, api.get(kind, ns, name))
                return self._send(200, {"kind": kind + "List", "items": api.list(kind, None if ns == "_all" else ns)})
            except ApiError as e:
                self._send(e.code, {"reason": e.reason, "message": str(e)})

        def do_POST(self):
            try:
                raw = self._body()
                if self.path == "/apply":
                    docs = [d for d in yaml.safe_load_all(raw.decode()) if d]
                    if len(docs) == 1 and isinstance(docs[0], list):
                        docs = docs[0]
                    out = []
                    for d in docs:
                        if not api.scheme.knows(d.get("kind", "")):
                            out.append({"kind": d.get("kind"), "name": d.get("metadata", {}).get("name"), "result": "skipped (kind not served here)"})
                            continue
                        o = api.apply(d)
                        out.append({"kind": o["kind"], "name": o["metadata"]["name"], "namespace": o["metadata"]["namespace"], "result": "applied"})
                    return self._send(200, out)
                r = self._route()
                if not r or not r[0]:
                    return self._send(404, {"reason": "NotFound", "message": self.path})
                obj = json.loads(raw or b"{}")
                obj.setdefault("kind", r[0])
                obj.setdefault("metadata", {}).setdefault("namespace", r[1])
                return self._send(201, api.create(obj))
            except ApiError as e:
                self._send(e.code, {"reason": e.reason, "message": str(e)})
            except Exception as e:
                self._send(400, {"reason": "BadRequest", "message": str(e)})

        def do_PUT(self):
            try:
                r = self._route()
                if not r or not r[0] or not r[2]:
                    return self._send(404, {"reason": "NotFound", "message": self.path})
                obj = json.loads(self._body() or b"{}")
                obj.setdefault("kind", r[0])
                return self._send(200, api.update_status(obj) if r[3] else api.update(obj))
            except ApiError as e:
                self._send(e.code, {"reason": e.reason, "message": str(e)})

        def do_DELETE(self):
            try:
                r = self._route()
                if not r or not r[0] or not r[2]:
                    return self._send(404, {"reason": "NotFound", "message": self.path})
                api.delete(r[0], r[1], r[2])
                return self._send(200, {"status": "Success"})
            except ApiError as e:
                self._send(e.code, {"reason": e.reason, "message": str(e)})

    return Handler


def serve(api: APIServer, host: str = "127.0.0.1", port: int = 8001, log_reader=None) -> ThreadingHTTPServer:
    srv = ThreadingHTTPServer((host, port), make_handler(api, log_reader))
    threading.Thread(target=srv.serve_forever, name="kubedl-rest", daemon=True).start()
    return srv
