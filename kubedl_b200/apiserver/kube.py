"""Client for a Kubernetes API server with the same interface as the in-process `APIServer`.

With it the operator runs the way the reference does — as a Deployment inside a cluster, or from
a laptop with a kubeconfig — against a real kube-apiserver: `Manager(api=KubeAPIServer(cfg))`.
The controllers, the engine and the informer cache are unchanged; only the object that answers
`create / get / list / update / update_status / delete / watch` differs.

  * credentials: `KUBECONFIG` (current context) first, else the in-cluster service account —
    the order of the reference (pkg/util/k8sutil/k8sutil.go:44-69); an explicit `--master` URL
    or `--kubeconfig` path wins over both.
  * reads and writes: plain REST, JSON bodies, `Status` errors mapped to the `ApiError` family.
  * `watch(kind, handler)`: one reflector thread per kind — LIST, then WATCH from the list's
    resourceVersion, re-WATCH from the last seen version when the stream ends, re-LIST on
    `410 Gone` (client-go's Reflector, which controller-runtime builds its cache on:
    reference main.go:70-75).  Handlers get `(type, obj, old)` like the in-process store's.
  * discovery: `crd_installed(kind)` asks `/apis/<group>/<version>` (reference
    pkg/util/workloadgate/workload_gate.go:90-106).
  * events: `record_event` POSTs a core/v1 Event (controller-runtime's recorder).

Only the standard library is used (http.client + ssl), so the operator image needs nothing extra.
"""
from __future__ import annotations

import base64
import collections
import copy  # noqa: F401
import http.client
import json
import os
import socket
import ssl
import tempfile
import threading
import time
import urllib.parse
import uuid
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Tuple

import yaml

from ..api import common as c
from ..api.common import deepcopy as _copy
from ..api.kinds import SCHEME
from ..utils.logger import logger
from . import store as st
from .store import ADDED, DELETED, MODIFIED, ApiError, WatchHandler

log = logger("kubedl.kube")

SERVICE_ACCOUNT_DIR = "/var/run/secrets/kubernetes.io/serviceaccount"
KUBECONFIG_ENV = "KUBECONFIG"


class ConfigError(Exception):
    pass


@dataclass
class KubeConfig:
    server: str                                  # https://host:port (or http:// for the façade)
    token: str = ""
    token_file: str = ""                         # re-read on every request (bound tokens rotate)
    ca_file: str = ""
    client_cert_file: str = ""
    client_key_file: str = ""
    insecure_skip_tls_verify: bool = False
    namespace: str = "default"
    username: str = ""
    password: str = ""
    _tmp: List[str] = field(default_factory=list, repr=False)

    # ------------------------------------------------------------------ loaders
    @classmethod
    def from_url(cls, url: str, token: str = "") -> "KubeConfig":
        if "://" not in url:
            url = "http://" + url
        return cls(server=url.rstrip("/"), token=token)

    @classmethod
    def in_cluster(cls, sa_dir: str = SERVICE_ACCOUNT_DIR, env=os.environ) -> "KubeConfig":
        host, port = env.get("KUBERNETES_SERVICE_HOST", ""), env.get("KUBERNETES_SERVICE_PORT", "") or "443"
        if not host:
            raise ConfigError("not running in a cluster: KUBERNETES_SERVICE_HOST is not set")
        token_file = os.path.join(sa_dir, "token")
        if not os.path.exists(token_file):
            raise ConfigError(f"service account token {token_file} not found")
        ns_file = os.path.join(sa_dir, "namespace")
        ns = open(ns_file).read().strip() if os.path.exists(ns_file) else "default"
        ca = os.path.join(sa_dir, "ca.crt")
        if ":" in host and not host.startswith("["):
            host = f"[{host}]"
        return cls(server=f"https://{host}:{port}", token_file=token_file,
                   ca_file=ca if os.path.exists(ca) else "", namespace=ns)

    @classmethod
    def from_kubeconfig(cls, path: str, context: Optional[str] = None) -> "KubeConfig":
        # KUBECONFIG may be a colon-separated list; the first file that has the context wins
        docs = []
        for p in [q for q in path.split(os.pathsep) if q]:
            p = os.path.expanduser(p)
            if not os.path.exists(p):
                continue
            with open(p) as f:
                d = yaml.safe_load(f) or {}
            d["_dir"] = os.path.dirname(os.path.abspath(p))
            docs.append(d)
        if not docs:
            raise ConfigError(f"kubeconfig {path!r} not found")
        ctx_name = context or next((d.get("current-context") for d in docs if d.get("current-context")), None)

        def find(section: str, name: str, key: str):
            for d in docs:
                for e in d.get(section) or []:
                    if e.get("name") == name:
                        return e.get(key) or {}, d["_dir"]
            raise ConfigError(f"kubeconfig has no {section[:-1]} named {name!r}")

        if not ctx_name:
            raise ConfigError("kubeconfig has no current-context")
        ctx, _ = find("contexts", ctx_name, "context")
        cluster, cdir = find("clusters", ctx.get("cluster", ""), "cluster")
        user, udir = find("users", ctx["user"], "user") if ctx.get("user") else ({}, "")
        if not cluster.get("server"):
            raise ConfigError(f"cluster {ctx.get('cluster')!r} has no server")
        cfg = cls(server=cluster["server"].rstrip("/"), namespace=ctx.get("namespace") or "default",
                  insecure_skip_tls_verify=bool(cluster.get("insecure-skip-tls-verify")))

        def material(d: dict, base: str, key: str) -> str:
            """`<key>-data` (base64, inline) or `<key>` (path, relative to the kubeconfig)."""
            data = d.get(key + "-data")
            if data:
                fd, name = tempfile.mkstemp(prefix="kubedl-kc-")
                with os.fdopen(fd, "wb") as f:
                    f.write(base64.b64decode(data))
                os.chmod(name, 0o600)
                cfg._tmp.append(name)
                return name
            p = d.get(key) or ""
            return p if not p or os.path.isabs(p) else os.path.join(base, p)

        cfg.ca_file = material(cluster, cdir, "certificate-authority")
        cfg.client_cert_file = material(user, udir, "client-certificate")
        cfg.client_key_file = material(user, udir, "client-key")
        cfg.token = user.get("token") or ""
        tf = user.get("tokenFile") or ""
        cfg.token_file = tf if not tf or os.path.isabs(tf) else os.path.join(udir, tf)
        cfg.username, cfg.password = user.get("username") or "", user.get("password") or ""
        if user.get("exec") or user.get("auth-provider"):
            raise ConfigError("kubeconfig users with exec / auth-provider plugins are not supported; "
                              "use a token, a token file or a client certificate")
        return cfg

    @classmethod
    def load(cls, master: str = "", kubeconfig: str = "", context: Optional[str] = None,
             env=os.environ, token_file: str = "") -> "KubeConfig":
        """Explicit flags first; then $KUBECONFIG; then the pod's service account."""
        if master and token_file and not kubeconfig:
            cfg = cls.from_url(master)
            cfg.token_file = token_file
            return cfg
        if kubeconfig:
            cfg = cls.from_kubeconfig(kubeconfig, context)
            if master:
                cfg.server = cls.from_url(master).server
            return cfg
        if master:
            return cls.from_url(master)
        if env.get(KUBECONFIG_ENV):
            return cls.from_kubeconfig(env[KUBECONFIG_ENV], context)
        return cls.in_cluster(env=env)

    # ------------------------------------------------------------------ use
    def bearer(self) -> str:
        if self.token_file:
            try:
                with open(self.token_file) as f:
                    return f.read().strip()
            except OSError:
                pass
        return self.token

    def headers(self) -> Dict[str, str]:
        h = {"Accept": "application/json", "User-Agent": "kubedl-b200"}
        tok = self.bearer()
        if tok:
            h["Authorization"] = "Bearer " + tok
        elif self.username:
            h["Authorization"] = "Basic " + base64.b64encode(f"{self.username}:{self.password}".encode()).decode()
        return h

    def ssl_context(self) -> Optional[ssl.SSLContext]:
        if not self.server.startswith("https://"):
            return None
        ctx = ssl.create_default_context(cafile=self.ca_file or None)
        if self.insecure_skip_tls_verify:
            ctx.check_hostname = False
            ctx.verify_mode = ssl.CERT_NONE
        if self.client_cert_file:
            ctx.load_cert_chain(self.client_cert_file, self.client_key_file or None)
        return ctx

    def connect(self, timeout: Optional[float]) -> http.client.HTTPConnection:
        u = urllib.parse.urlsplit(self.server)
        if u.scheme == "https":
            conn = http.client.HTTPSConnection(u.hostname, u.port or 443, timeout=timeout, context=self.ssl_context())
        else:
            conn = http.client.HTTPConnection(u.hostname, u.port or 80, timeout=timeout)
        conn.connect()
        conn.sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)   # small requests, latency matters
        return conn

    @property
    def base_path(self) -> str:
        return urllib.parse.urlsplit(self.server).path.rstrip("/")

    def cleanup(self) -> None:
        for p in self._tmp:
            try:
                os.unlink(p)
            except OSError:
                pass
        self._tmp.clear()


_ERRORS = {"NotFound": st.NotFound, "AlreadyExists": st.AlreadyExists, "Conflict": st.Conflict,
           "Invalid": st.Invalid, "Timeout": st.Timeout, "ServerTimeout": st.Timeout, "Expired": st.Gone,
           "Gone": st.Gone}


class Forbidden(ApiError):
    reason, code = "Forbidden", 403


class Unauthorized(ApiError):
    reason, code = "Unauthorized", 401


_ERRORS.update({"Forbidden": Forbidden, "Unauthorized": Unauthorized})


def error_from_status(code: int, body: bytes) -> ApiError:
    """HTTP status + `Status` body -> the matching ApiError subclass (apierrors.IsNotFound etc.)."""
    reason, message = "", body.decode(errors="replace")[:500]
    try:
        d = json.loads(body)
        if isinstance(d, dict):
            reason, message = d.get("reason") or "", d.get("message") or message
            code = int(d.get("code") or code)
    except ValueError:
        pass
    cls = _ERRORS.get(reason)
    if cls is None:
        cls = {404: st.NotFound, 409: st.Conflict, 410: st.Gone, 422: st.Invalid, 504: st.Timeout,
               401: Unauthorized, 403: Forbidden}.get(code)
    if cls is None:
        err = ApiError(message)
        err.code, err.reason = code, reason or "InternalError"
        return err
    return cls(message)


class _Reflector:
    """LIST + WATCH of one kind, fanned out to handlers (client-go Reflector + shared informer)."""

    def __init__(self, client: "KubeAPIServer", kind: str):
        self.client, self.kind = client, kind
        self.handlers: List[WatchHandler] = []
        self.cache: Dict[Tuple[str, str], dict] = {}
        self.lock = threading.RLock()
        self.synced = threading.Event()
        self.stopping = threading.Event()
        self.relists = 0
        self.shared: set = set()
        self._conn: Optional[http.client.HTTPConnection] = None
        self.thread = threading.Thread(target=self._run, name=f"reflector-{kind}", daemon=True)
        self.thread.start()

    # -- handler management
    def add(self, handler: WatchHandler, replay: bool, shared: bool = False) -> Callable[[], None]:
        if not self.synced.wait(self.client.sync_timeout):
            raise st.Timeout(f"timed out waiting for the {self.kind} cache to sync")
        with self.lock:
            if replay:
                for o in list(self.cache.values()):
                    handler(ADDED, _copy(o), None)
            self.handlers.append(handler)
            if shared:
                self.shared.add(handler)

        def stop():
            with self.lock:
                if handler in self.handlers:
                    self.handlers.remove(handler)
                self.shared.discard(handler)
        return stop

    def _emit(self, ev: str, obj: dict) -> None:
        key = (c.namespace_of(obj), c.name_of(obj))
        with self.lock:
            old = self.cache.get(key)
            if ev == DELETED:
                if old is None:
                    return
                self.cache.pop(key, None)
                old = None
            else:
                if old is not None and old["metadata"].get("resourceVersion") == obj["metadata"].get("resourceVersion"):
                    return                      # replayed by a re-list / re-watch, nothing new
                ev = MODIFIED if old is not None else ADDED
                self.cache[key] = obj
            snap = None
            for h in list(self.handlers):
                try:
                    if h in self.shared:        # read-only handlers share one snapshot (store.APIServer.watch)
                        if snap is None:
                            snap = (_copy(obj), _copy(old) if old is not None else None)
                            if st._FREEZE_SNAPSHOTS:
                                snap = (st._freeze(snap[0]), st._freeze(snap[1]) if snap[1] is not None else None)
                        h(ev, snap[0], snap[1])
                    else:
                        h(ev, _copy(obj), _copy(old) if old is not None else None)
                except Exception as e:          # a handler must not take the reflector down
                    log.error("watch handler for %s failed: %r", self.kind, e)

    # -- the loop
    def _relist(self) -> str:
        items, rv = self.client.list_with_rv(self.kind)
        self.relists += 1
        seen = set()
        for o in items:
            o.setdefault("kind", self.kind)
            seen.add((c.namespace_of(o), c.name_of(o)))
            self._emit(MODIFIED, o)
        with self.lock:
            gone = [o for k, o in self.cache.items() if k not in seen]
        for o in gone:                          # deleted while we were not watching
            self._emit(DELETED, o)
        self.synced.set()
        return rv

    def _watch_once(self, rv: str) -> str:
        g = self.client
        path = g._collection_path(self.kind, None) + "?" + urllib.parse.urlencode(
            {"watch": "true", "resourceVersion": rv, "timeoutSeconds": str(int(g.watch_timeout)),
             "allowWatchBookmarks": "true"})
        conn = g.config.connect(timeout=g.watch_timeout + 15)
        self._conn = conn
        try:
            conn.request("GET", g.config.base_path + path, headers=g.config.headers())
            resp = conn.getresponse()
            if resp.status != 200:
                raise error_from_status(resp.status, resp.read())
            while not self.stopping.is_set():
                line = resp.readline()
                if not line:
                    return rv                   # server closed the stream (timeoutSeconds): watch again
                line = line.strip()
                if not line:
                    continue
                evt = json.loads(line)
                typ, obj = evt.get("type"), evt.get("object") or {}
                if typ == "ERROR":
                    raise error_from_status(int(obj.get("code") or 500), json.dumps(obj).encode())
                new_rv = (obj.get("metadata") or {}).get("resourceVersion")
                if typ in (ADDED, MODIFIED, DELETED):
                    obj.setdefault("kind", self.kind)
                    self._emit(typ, obj)
                if new_rv:
                    rv = new_rv                 # BOOKMARK events only move the version forward
            return rv
        finally:
            self._conn = None
            try:
                conn.close()
            except Exception:
                pass

    def _run(self) -> None:
        backoff = 0.05
        while not self.stopping.is_set():
            try:
                rv = self._relist()
                backoff = 0.05
                while not self.stopping.is_set():
                    rv = self._watch_once(rv)
            except st.Gone:
                continue                        # history expired: list again
            except Exception as e:
                if self.stopping.is_set():
                    break
                log.warning("reflector %s: %r; retrying in %.2fs", self.kind, e, backoff)
                self.stopping.wait(backoff)
                backoff = min(backoff * 2, 5.0)

    def stop(self) -> None:
        self.stopping.set()
        conn = self._conn
        if conn is not None and conn.sock is not None:
            try:
                conn.sock.shutdown(2)           # unblocks the readline() of the watch
            except OSError:
                pass


def _url_template(path: str) -> str:
    """`/api/v1/namespaces/kubedl/pods/x-worker-0?y` -> `/api/v1/namespaces/{namespace}/pods/{name}`
    (client-go labels request latency by URL template, not by object)."""
    parts = path.split("?")[0].strip("/").split("/")
    out, i = [], 0
    head = 2 if parts and parts[0] == "api" else 3
    out += parts[:head]
    rest_ = parts[head:]
    if len(rest_) >= 2 and rest_[0] == "namespaces" and len(rest_) > 2:
        out += ["namespaces", "{namespace}"]
        rest_ = rest_[2:]
    if rest_:
        out.append(rest_[0])
        if len(rest_) > 1:
            out.append("{name}")
        out += rest_[2:]
    return "/" + "/".join(out)


class KubeAPIServer:
    """Drop-in for `store.APIServer` backed by a remote Kubernetes API server."""

    def __init__(self, config: KubeConfig, scheme=SCHEME, request_timeout: float = 30.0,
                 watch_timeout: float = 240.0, sync_timeout: float = 60.0):
        self.config, self.scheme = config, scheme
        self.request_timeout, self.watch_timeout, self.sync_timeout = request_timeout, watch_timeout, sync_timeout
        self._local = threading.local()
        self._reflectors: Dict[str, _Reflector] = {}
        self._lock = threading.RLock()
        self._discovery: Dict[str, Tuple[float, List[str]]] = {}
        self.discovery_ttl = 10.0
        self.request_log = collections.deque(maxlen=65536)
        self.remote = True
        self.metrics = None      # plugins/controller_metrics.ControllerMetrics (rest_client_* series)
        self._host = urllib.parse.urlsplit(config.server).netloc

    # ------------------------------------------------------------------ paths
    def _prefix(self, kind: str) -> Tuple[str, str]:
        try:
            g, v, plural = self.scheme.resource(kind)
        except KeyError:
            raise st.NotFound(f"the server could not find the requested resource ({kind})")
        return (f"/apis/{g}/{v}" if g else f"/api/{v}"), plural

    def _collection_path(self, kind: str, namespace: Optional[str]) -> str:
        prefix, plural = self._prefix(kind)
        if namespace and kind not in ("Namespace", "Node"):      # cluster-scoped kinds have no namespace segment
            return f"{prefix}/namespaces/{urllib.parse.quote(namespace)}/{plural}"
        return f"{prefix}/{plural}"

    def _object_path(self, kind: str, namespace: str, name: str, sub: str = "") -> str:
        return self._collection_path(kind, namespace or "default") + "/" + urllib.parse.quote(name) + (f"/{sub}" if sub else "")

    # ------------------------------------------------------------------ transport
    def _request(self, method: str, path: str, body: Optional[dict] = None,
                 content_type: str = "application/json") -> dict:
        data = json.dumps(body).encode() if body is not None else None
        headers = self.config.headers()
        if data is not None:
            headers["Content-Type"] = content_type
        last: Optional[Exception] = None
        for attempt in range(2):                 # a kept-alive connection may have been closed under us
            conn = getattr(self._local, "conn", None)
            t0 = time.perf_counter()
            try:
                if conn is None:
                    conn = self._local.conn = self.config.connect(self.request_timeout)
                conn.request(method, self.config.base_path + path, body=data, headers=headers)
                resp = conn.getresponse()
                raw = resp.read()
            except (http.client.HTTPException, ConnectionError, BrokenPipeError, TimeoutError, OSError) as e:
                last = e
                try:
                    if conn is not None:
                        conn.close()
                finally:
                    self._local.conn = None
                continue
            if self.metrics is not None:
                self.metrics.observe_request(method, self._host, _url_template(path), resp.status, time.perf_counter() - t0)
            if resp.status >= 400:
                raise error_from_status(resp.status, raw)
            if not raw:
                return {}
            try:
                return json.loads(raw)
            except ValueError:
                return {"raw": raw.decode(errors="replace")}
        raise st.Timeout(f"{method} {path}: {last!r}")

    def _prepare(self, obj: dict) -> dict:
        kind = obj.get("kind")
        if not kind or not c.name_of(obj):
            raise st.Invalid("object needs kind and metadata.name")
        o = _copy(obj)
        o.setdefault("apiVersion", self.scheme.api_version(kind) if self.scheme.knows(kind) else "v1")
        o.setdefault("metadata", {})["namespace"] = c.namespace_of(obj)
        return o

    def _stamp(self, kind: str, obj: dict) -> dict:
        obj.setdefault("kind", kind)             # items of a List come without kind / apiVersion
        if self.scheme.knows(kind):
            obj.setdefault("apiVersion", self.scheme.api_version(kind))
        return obj

    # ------------------------------------------------------------------ discovery
    def server_resources_for_group_version(self, gv: str) -> List[str]:
        now = time.monotonic()
        hit = self._discovery.get(gv)
        if hit and now - hit[0] < self.discovery_ttl:
            return hit[1]
        try:
            d = self._request("GET", ("/apis/" if "/" in gv else "/api/") + gv)
            kinds = sorted({r["kind"] for r in d.get("resources", []) if "/" not in r.get("name", "")})
        except st.NotFound:
            kinds = []
        self._discovery[gv] = (now, kinds)
        return kinds

    def crd_installed(self, kind: str) -> bool:
        if not self.scheme.knows(kind):
            return False
        return kind in self.server_resources_for_group_version(self.scheme.api_version(kind))

    def install_crd(self, kind: str) -> None:
        """Create the CustomResourceDefinition from `api/crd.py` if the server lacks it."""
        from ..api import crd as crd_mod
        if self.crd_installed(kind):
            return
        info = self.scheme.info(kind)
        for doc in (crd_mod.crd_for_v1(info), crd_mod.crd_for(info)):   # v1 first, v1beta1 for old clusters
            try:
                self._request("POST", f"/apis/{doc['apiVersion']}/customresourcedefinitions", doc)
                break
            except st.AlreadyExists:
                break
            except st.NotFound:
                continue
        self._discovery.clear()

    # ------------------------------------------------------------------ verbs
    def create(self, obj: dict) -> dict:
        o = self._prepare(obj)
        kind, ns = o["kind"], o["metadata"]["namespace"]
        self.request_log.append(("create", kind, f"{ns}/{c.name_of(o)}"))
        return self._stamp(kind, self._request("POST", self._collection_path(kind, ns), o))

    def get(self, kind: str, namespace: str, name: str) -> dict:
        return self._stamp(kind, self._request("GET", self._object_path(kind, namespace, name)))

    def list_with_rv(self, kind: str, namespace: Optional[str] = None,
                     label_selector: Optional[Dict[str, str]] = None) -> Tuple[List[dict], str]:
        path = self._collection_path(kind, namespace)
        if label_selector:
            sel = ",".join(f"{k}={v}" for k, v in sorted(label_selector.items()))
            path += "?" + urllib.parse.urlencode({"labelSelector": sel})
        d = self._request("GET", path)
        items = [self._stamp(kind, o) for o in d.get("items") or []]
        items.sort(key=lambda o: (c.namespace_of(o), c.name_of(o)))
        return items, str((d.get("metadata") or {}).get("resourceVersion") or "0")

    def list(self, kind: str, namespace: Optional[str] = None,
             label_selector: Optional[Dict[str, str]] = None) -> List[dict]:
        return self.list_with_rv(kind, namespace, label_selector)[0]

    def update(self, obj: dict) -> dict:
        o = self._prepare(obj)
        kind, ns, name = o["kind"], o["metadata"]["namespace"], c.name_of(o)
        self.request_log.append(("update", kind, f"{ns}/{name}"))
        return self._stamp(kind, self._request("PUT", self._object_path(kind, ns, name), o))

    def update_status(self, obj: dict) -> dict:
        o = self._prepare(obj)
        kind, ns, name = o["kind"], o["metadata"]["namespace"], c.name_of(o)
        self.request_log.append(("update_status", kind, f"{ns}/{name}"))
        return self._stamp(kind, self._request("PUT", self._object_path(kind, ns, name, "status"), o))

    def patch_labels(self, kind: str, namespace: str, name: str, labels: Dict[str, Optional[str]]) -> dict:
        return self._stamp(kind, self._request("PATCH", self._object_path(kind, namespace, name),
                                               {"metadata": {"labels": dict(labels)}},
                                               content_type="application/merge-patch+json"))

    def merge_patch(self, kind: str, namespace: str, name: str, patch: dict) -> dict:
        self.request_log.append(("patch", kind, f"{namespace or 'default'}/{name}"))
        return self._stamp(kind, self._request("PATCH", self._object_path(kind, namespace, name), patch,
                                               content_type="application/merge-patch+json"))

    def delete(self, kind: str, namespace: str, name: str, cascade: bool = True) -> None:
        self.request_log.append(("delete", kind, f"{namespace or 'default'}/{name}"))
        policy = "Background" if cascade else "Orphan"
        self._request("DELETE", self._object_path(kind, namespace, name) + f"?propagationPolicy={policy}",
                      {"kind": "DeleteOptions", "apiVersion": "v1", "propagationPolicy": policy})

    def apply(self, obj: dict) -> dict:
        try:
            return self.create(obj)
        except st.AlreadyExists:
            cur = self.get(obj["kind"], c.namespace_of(obj), c.name_of(obj))
            merged = _copy(obj)
            merged.setdefault("metadata", {})["resourceVersion"] = cur["metadata"].get("resourceVersion")
            return self.update(merged)

    # ------------------------------------------------------------------ watch
    def watch(self, kind: str, handler: WatchHandler, replay: bool = True, shared: bool = False) -> Callable[[], None]:
        with self._lock:
            r = self._reflectors.get(kind)
            if r is None:
                r = self._reflectors[kind] = _Reflector(self, kind)
        return r.add(handler, replay, shared)

    def reflector(self, kind: str) -> Optional[_Reflector]:
        return self._reflectors.get(kind)

    # ------------------------------------------------------------------ events
    def record_event(self, involved: dict, etype: str, reason: str, message: str, source: str = "kubedl") -> dict:
        ns = c.namespace_of(involved)
        now = c.now_rfc3339()
        ev = {"apiVersion": "v1", "kind": "Event",
              "metadata": {"name": f"{c.name_of(involved)}.{uuid.uuid4().hex[:12]}", "namespace": ns},
              "involvedObject": {"kind": involved.get("kind", ""), "namespace": ns, "name": c.name_of(involved),
                                 "uid": c.uid_of(involved), "apiVersion": involved.get("apiVersion", "")},
              "type": etype, "reason": reason, "message": message, "count": 1,
              "firstTimestamp": now, "lastTimestamp": now, "source": {"component": source}}
        return self.create(ev)

    def events_for(self, involved: dict) -> List[dict]:
        uid = c.uid_of(involved)
        return [e for e in self.list("Event", c.namespace_of(involved))
                if e.get("involvedObject", {}).get("uid") == uid]

    # ------------------------------------------------------------------ lifecycle
    def healthy(self) -> bool:
        try:
            self._request("GET", "/healthz")
            return True
        except Exception:
            return False

    def close(self) -> None:
        with self._lock:
            for r in self._reflectors.values():
                r.stop()
            self._reflectors.clear()
        self.config.cleanup()
