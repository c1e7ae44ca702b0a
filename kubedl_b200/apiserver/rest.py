"""REST front of the in-process API server, speaking the Kubernetes wire protocol.

`kubedlctl`, curl and — most importantly — the operator itself (through
`apiserver/kube.py::KubeAPIServer`, the client it also uses against a real kube-apiserver) talk
to it.  The reference reaches its API server through client-go / controller-runtime
(reference main.go:70-75, pkg/util/k8sutil/k8sutil.go:44-69); this is the server side of that
conversation for a single box, and the test double for the client.

  GET  /api/v1, /apis, /apis/<group>/<version>                          discovery
  GET/POST   /apis/<group>/<version>/namespaces/<ns>/<plural>          list / create
  GET  /apis/<group>/<version>/<plural>, /api/v1/<plural>              list, all namespaces
  GET/PUT/PATCH/DELETE  .../<plural>/<name>[/status]                   get / replace / merge-patch / delete
  GET  ...?watch=true&resourceVersion=N[&timeoutSeconds=S]             chunked stream of watch events
  GET  ...?labelSelector=a=b,c=d                                       equality selectors
  POST /apply            multi-document YAML or JSON list              create-or-replace
  GET  /logs/<ns>/<pod>                                                pod log from the node agent
  GET  /healthz, /readyz, /version
Errors are `Status` objects (`reason`, `code`, `message`), as client-go expects.
"""
from __future__ import annotations

import json
import queue
import socket
import threading
import time
import urllib.parse
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Callable, Dict, List, Optional, Tuple

import yaml

from ..api import common as c
from ..api.kinds import SCHEME
from .store import ADDED, DELETED, ApiError, APIServer, Gone


def _resource_map(scheme) -> Dict[Tuple[str, str, str], str]:
    """(group, version, plural) -> kind for every kind the scheme knows."""
    return {scheme.resource(k): k for k in scheme.all_kinds()}


def parse_label_selector(text: Optional[str]) -> Optional[Dict[str, str]]:
    """`a=b,c==d` -> {a: b, c: d}.  Only equality terms (all the operator ever sends)."""
    if not text:
        return None
    out = {}
    for term in text.split(","):
        term = term.strip()
        if not term:
            continue
        if "!=" in term or " in " in term or " notin " in term:
            raise ValueError(f"unsupported selector term {term!r}")
        k, _, v = term.partition("==" if "==" in term else "=")
        out[k.strip()] = v.strip()
    return out


def merge_patch(target, patch):
    """RFC 7386 JSON merge patch."""
    if not isinstance(patch, dict):
        return patch
    if not isinstance(target, dict):
        target = {}
    out = dict(target)
    for k, v in patch.items():
        if v is None:
            out.pop(k, None)
        else:
            out[k] = merge_patch(out.get(k), v)
    return out


def url_tail(path: str) -> str:
    return path.split("?")[0].rstrip("/").rsplit("/", 1)[-1]


def status_body(code: int, reason: str, message: str) -> dict:
    return {"kind": "Status", "apiVersion": "v1", "metadata": {}, "status": "Failure" if code >= 400 else "Success",
            "message": message, "reason": reason, "code": code}


class Route:
    __slots__ = ("kind", "namespace", "name", "status", "query")

    def __init__(self, kind, namespace, name, status, query):
        self.kind, self.namespace, self.name, self.status, self.query = kind, namespace, name, status, query

    def q(self, key: str, default: Optional[str] = None) -> Optional[str]:
        v = self.query.get(key)
        return v[0] if v else default


def make_handler(api: APIServer, log_reader: Optional[Callable[[str, str], str]] = None,
                 stopping: Optional[threading.Event] = None, latency: float = 0.0, token: str = ""):
    """latency: seconds added to every non-watch request — a stand-in for the network round trip
    and etcd commit of a real API server when measuring what the operator's request pattern costs."""
    resources = _resource_map(api.scheme)
    stopping = stopping or threading.Event()

    def served(kind: str) -> bool:
        return not api.scheme.is_workload(kind) or api.crd_installed(kind)

    def resource_list(group: str, version: str) -> dict:
        out = []
        for (g, v, plural), kind in sorted(resources.items()):
            if (g, v) != (group, version) or not served(kind):
                continue
            verbs = ["create", "delete", "get", "list", "patch", "update", "watch"]
            out.append({"name": plural, "singularName": kind.lower(), "namespaced": kind != "Namespace",
                        "kind": kind, "verbs": verbs})
            if api.scheme.is_workload(kind) or kind == "Pod":
                out.append({"name": plural + "/status", "singularName": "", "namespaced": True, "kind": kind,
                            "verbs": ["get", "patch", "update"]})
        return {"kind": "APIResourceList", "apiVersion": "v1",
                "groupVersion": f"{group}/{version}" if group else version, "resources": out}

    def group_list() -> dict:
        groups: Dict[str, List[str]] = {}
        for (g, v, _), kind in resources.items():
            if g and served(kind) and v not in groups.setdefault(g, []):
                groups[g].append(v)
        return {"kind": "APIGroupList", "apiVersion": "v1",
                "groups": [{"name": g, "versions": [{"groupVersion": f"{g}/{v}", "version": v} for v in vs],
                            "preferredVersion": {"groupVersion": f"{g}/{vs[0]}", "version": vs[0]}}
                           for g, vs in sorted(groups.items())],
                # kept for older kubedlctl builds
                "kinds": [k for k in api.scheme.kinds() if api.crd_installed(k)]}

    def crd_request(method: str, path: str, body: bytes):
        """CustomResourceDefinitions: creating one turns the kind on, deleting it turns it off.
        Only kinds the scheme has code for can be served.  Returns (code, body) or None."""
        parts = [p for p in path.split("/") if p]
        if len(parts) < 4 or parts[:2] != ["apis", "apiextensions.k8s.io"] or parts[3] != "customresourcedefinitions":
            return None
        from ..api import crd as crd_mod
        by_name = {i.crd_name: i for i in api.scheme.workload_kinds()}
        render = crd_mod.crd_for_v1 if parts[2] == "v1" else crd_mod.crd_for
        if method == "POST":
            doc = json.loads(body or b"{}")
            info = by_name.get((doc.get("metadata") or {}).get("name", ""))
            if info is None:
                return 422, status_body(422, "Invalid", "no controller in this build serves that CustomResourceDefinition")
            if api.crd_installed(info.kind):
                return 409, status_body(409, "AlreadyExists", f'customresourcedefinitions "{info.crd_name}" already exists')
            api.install_crd(info.kind)
            return 201, render(info)
        name = parts[4] if len(parts) > 4 else None
        if method == "GET" and name is None:
            return 200, {"kind": "CustomResourceDefinitionList", "apiVersion": f"apiextensions.k8s.io/{parts[2]}",
                         "items": [render(i) for i in by_name.values() if api.crd_installed(i.kind)]}
        info = by_name.get(name or "")
        if info is None or not api.crd_installed(info.kind):
            return 404, status_body(404, "NotFound", f'customresourcedefinitions "{name}" not found')
        if method == "GET":
            return 200, render(info)
        if method == "DELETE":
            api.uninstall_crd(info.kind)
            return 200, dict(status_body(200, "", ""), status="Success")
        return None

    class Handler(BaseHTTPRequestHandler):
        protocol_version = "HTTP/1.1"

        def log_message(self, *a):  # quiet
            pass

        def setup(self):
            super().setup()
            # headers and body leave in two writes; without this Nagle + delayed ACK cost 40 ms per reply
            self.request.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)

        # ------------------------------------------------------------ plumbing
        def _send(self, code: int, body) -> None:
            if latency > 0:
                time.sleep(latency)
            data = body if isinstance(body, bytes) else json.dumps(body).encode()
            self.send_response(code)
            self.send_header("Content-Type", "application/json" if not isinstance(body, bytes) else "text/plain")
            self.send_header("Content-Length", str(len(data)))
            self.end_headers()
            self.wfile.write(data)

        def _fail(self, e: Exception) -> None:
            if isinstance(e, ApiError):
                return self._send(e.code, status_body(e.code, e.reason, str(e)))
            self._send(400, status_body(400, "BadRequest", str(e)))

        def _not_found(self) -> None:
            self._send(404, status_body(404, "NotFound", f"the server could not find the requested resource ({self.path})"))

        def _authorized(self) -> bool:
            """Static bearer token (`--token-file`): enough to keep strangers out when the server has to
            listen beyond loopback for node agents on other boxes; TLS is the deployment's job (tunnel / proxy)."""
            if not token or self.path.split("?")[0] in ("/healthz", "/readyz", "/livez"):
                return True
            import hmac
            got = self.headers.get("Authorization", "")
            if hmac.compare_digest(got.encode(), f"Bearer {token}".encode()):
                return True
            self._body()
            self._send(401, status_body(401, "Unauthorized", "Unauthorized"))
            return False

        def _body(self) -> bytes:
            n = int(self.headers.get("Content-Length", "0") or 0)
            return self.rfile.read(n) if n else b""

        def _route(self) -> Optional[Route]:
            url = urllib.parse.urlsplit(self.path)
            query = urllib.parse.parse_qs(url.query)
            parts = [urllib.parse.unquote(p) for p in url.path.split("/") if p]
            if len(parts) >= 2 and parts[0] == "api":
                group, version, rest_ = "", parts[1], parts[2:]
            elif len(parts) >= 3 and parts[0] == "apis":
                group, version, rest_ = parts[1], parts[2], parts[3:]
            else:
                return None
            ns = None
            if len(rest_) >= 2 and rest_[0] == "namespaces" and not (group == "" and len(rest_) == 2):
                ns, rest_ = rest_[1], rest_[2:]
            if not rest_:
                return None
            kind = resources.get((group, version, rest_[0]))
            if kind is None or not served(kind):
                return None
            if ns == "_all":
                ns = None
            name = rest_[1] if len(rest_) > 1 else None
            return Route(kind, ns, name, len(rest_) > 2 and rest_[2] == "status", query)

        # ------------------------------------------------------------ watch stream
        def _chunk(self, payload: bytes) -> None:
            self.wfile.write(f"{len(payload):x}\r\n".encode() + payload + b"\r\n")
            self.wfile.flush()

        def _watch(self, r: Route) -> None:
            selector = parse_label_selector(r.q("labelSelector"))
            timeout = float(r.q("timeoutSeconds") or 300)
            rv = r.q("resourceVersion")
            events: "queue.Queue[Tuple[str, dict]]" = queue.Queue()
            # subscribe first, then replay: nothing can fall between the two
            # the stream only serialises what it is handed: one snapshot shared with the other read-only watchers
            stop = api.watch(r.kind, lambda ev, obj, old: events.put((ev, obj)), replay=False, shared=True)
            try:
                try:
                    if rv in (None, "", "0"):
                        items, list_rv = api.list_with_rv(r.kind)
                        backlog = [(ADDED, o) for o in items]
                        sent_rv = int(list_rv)
                    else:
                        backlog = api.events_since(r.kind, rv)
                        sent_rv = int(rv)
                except Gone as e:
                    # client-go expects 200 + an ERROR event carrying a 410 Status
                    self.send_response(200)
                    self.send_header("Content-Type", "application/json")
                    self.send_header("Transfer-Encoding", "chunked")
                    self.end_headers()
                    self._chunk(json.dumps({"type": "ERROR", "object": status_body(410, "Expired", str(e))}).encode() + b"\n")
                    self.wfile.write(b"0\r\n\r\n")
                    return
                self.send_response(200)
                self.send_header("Content-Type", "application/json")
                self.send_header("Transfer-Encoding", "chunked")
                self.end_headers()

                def wanted(obj: dict) -> bool:
                    if r.namespace and c.namespace_of(obj) != r.namespace:
                        return False
                    if r.name and c.name_of(obj) != r.name:
                        return False
                    return not selector or all(c.labels_of(obj).get(k) == v for k, v in selector.items())

                def emit(ev: str, obj: dict) -> None:
                    self._chunk(json.dumps({"type": ev, "object": obj}).encode() + b"\n")

                replayed_list = rv in (None, "", "0")
                for ev, obj in backlog:
                    orv = int(obj["metadata"].get("resourceVersion", 0))
                    if wanted(obj):
                        emit(ev, obj)
                    if not replayed_list:
                        sent_rv = max(sent_rv, orv)
                deadline = time.monotonic() + timeout
                while not stopping.is_set():
                    left = deadline - time.monotonic()
                    if left <= 0:
                        break
                    try:
                        ev, obj = events.get(timeout=min(left, 0.25))
                    except queue.Empty:
                        continue
                    orv = int(obj["metadata"].get("resourceVersion", 0))
                    if orv <= sent_rv:
                        continue      # already covered by the replay
                    sent_rv = orv
                    if wanted(obj):
                        emit(ev, obj)
                self.wfile.write(b"0\r\n\r\n")
                self.wfile.flush()
            except (BrokenPipeError, ConnectionResetError, OSError):
                self.close_connection = True
            finally:
                stop()

        # ------------------------------------------------------------ verbs
        def do_GET(self):
            if not self._authorized():
                return
            try:
                path = self.path.split("?")[0].rstrip("/")
                if path in ("/healthz", "/readyz", "/livez"):
                    return self._send(200, {"status": "ok"})
                if path == "/version":
                    return self._send(200, {"major": "1", "minor": "16", "gitVersion": "v1.16.0-kubedl-b200", "platform": "linux/amd64"})
                if path == "/apis":
                    return self._send(200, group_list())
                if path == "/api":
                    return self._send(200, {"kind": "APIVersions", "versions": ["v1"]})
                if path == "/api/v1":
                    return self._send(200, resource_list("", "v1"))
                parts = [p for p in path.split("/") if p]
                if len(parts) == 3 and parts[0] == "apis":
                    rl = resource_list(parts[1], parts[2])
                    return self._send(200, rl) if rl["resources"] else self._not_found()
                if path.startswith("/logs/"):
                    _, _, ns, pod = path.split("/", 3)
                    text = log_reader(ns, pod) if log_reader else ""
                    return self._send(200, text.encode())
                hit = crd_request("GET", path, b"")
                if hit:
                    return self._send(*hit)
                r = self._route()
                if r is None:
                    return self._not_found()
                if (r.q("watch") or "").lower() in ("1", "true"):
                    return self._watch(r)
                if r.name and r.kind == "Pod" and url_tail(self.path) == "log":
                    api.get("Pod", r.namespace or "default", r.name)          # 404 for unknown pods
                    text = log_reader(r.namespace or "default", r.name) if log_reader else ""
                    return self._send(200, text.encode())
                if r.name:
                    return self._send(200, api.get(r.kind, r.namespace or "default", r.name))
                items, rv = api.list_with_rv(r.kind, r.namespace, parse_label_selector(r.q("labelSelector")))
                return self._send(200, {"kind": r.kind + "List", "apiVersion": api.scheme.api_version(r.kind),
                                        "metadata": {"resourceVersion": rv}, "items": items})
            except Exception as e:
                self._fail(e)

        def do_POST(self):
            if not self._authorized():
                return
            try:
                raw = self._body()
                if self.path.split("?")[0] == "/apply":
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
                hit = crd_request("POST", self.path.split("?")[0], raw)
                if hit:
                    return self._send(*hit)
                r = self._route()
                if r is None or r.name:
                    return self._not_found()
                obj = json.loads(raw or b"{}")
                obj.setdefault("kind", r.kind)
                obj.setdefault("metadata", {}).setdefault("namespace", r.namespace or "default")
                return self._send(201, api.create(obj))
            except Exception as e:
                self._fail(e)

        def do_PUT(self):
            if not self._authorized():
                return
            try:
                r = self._route()
                if r is None or not r.name:
                    return self._not_found()
                obj = json.loads(self._body() or b"{}")
                obj.setdefault("kind", r.kind)
                obj.setdefault("metadata", {}).setdefault("namespace", r.namespace or "default")
                obj["metadata"].setdefault("name", r.name)
                return self._send(200, api.update_status(obj) if r.status else api.update(obj))
            except Exception as e:
                self._fail(e)

        def do_PATCH(self):
            if not self._authorized():
                return
            try:
                r = self._route()
                if r is None or not r.name:
                    return self._not_found()
                patch = json.loads(self._body() or b"{}")
                if not isinstance(patch, dict):
                    raise ValueError("only merge patches are supported")
                cur = api.get(r.kind, r.namespace or "default", r.name)
                new = merge_patch(cur, patch)
                if "resourceVersion" not in (patch.get("metadata") or {}):
                    new["metadata"].pop("resourceVersion", None)   # unconditional unless the patch pins it
                return self._send(200, api.update_status(new) if r.status else api.update(new))
            except Exception as e:
                self._fail(e)

        def do_DELETE(self):
            if not self._authorized():
                return
            try:
                hit = crd_request("DELETE", self.path.split("?")[0], b"")
                if hit:
                    self._body()
                    return self._send(*hit)
                r = self._route()
                if r is None or not r.name:
                    return self._not_found()
                policy = r.q("propagationPolicy")
                raw = self._body()
                if raw:
                    try:
                        policy = json.loads(raw).get("propagationPolicy", policy)
                    except ValueError:
                        pass
                api.delete(r.kind, r.namespace or "default", r.name, cascade=policy != "Orphan")
                return self._send(200, dict(status_body(200, "", ""), status="Success",
                                            details={"name": r.name, "kind": r.kind.lower() + "s"}))
            except Exception as e:
                self._fail(e)

    return Handler


class RestServer(ThreadingHTTPServer):
    daemon_threads = True
    allow_reuse_address = True

    def __init__(self, addr, api: APIServer, log_reader=None, latency: float = 0.0, token: str = ""):
        self.stopping = threading.Event()
        super().__init__(addr, make_handler(api, log_reader, self.stopping, latency, token))

    @property
    def url(self) -> str:
        host, port = self.server_address[:2]
        return f"http://{host}:{port}"

    def shutdown(self) -> None:
        self.stopping.set()          # ends the open watch streams
        super().shutdown()
        self.server_close()


def serve(api: APIServer, host: str = "127.0.0.1", port: int = 8001, log_reader=None, latency: float = 0.0,
          token: str = "") -> RestServer:
    srv = RestServer((host, port), api, log_reader, latency, token)
    threading.Thread(target=srv.serve_forever, name="kubedl-rest", daemon=True).start()
    return srv
