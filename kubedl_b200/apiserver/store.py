"""In-process API server: the cluster state the control plane reconciles against.

The reference talks to a real kube-apiserver through controller-runtime (main.go:70-75) and its
tests use controller-runtime's fake client (api/pytorch/v1/v1_suite_test.go:31-41).  This store
is the B200-box equivalent of both: typed-by-kind object storage with uid / resourceVersion /
creationTimestamp, optimistic concurrency, the status subresource, label-selector LIST, WATCH
streams (ADDED / MODIFIED / DELETED), ownerReference garbage collection, CRD discovery and
Events.  Objects are Kubernetes wire-format dicts.
"""
from __future__ import annotations

import collections
import copy  # noqa: F401
import logging
import threading
import traceback
import uuid
from typing import Callable, Deque, Dict, List, Optional, Tuple

from ..api import common as c
from ..api.common import deepcopy as _copy
from ..api.kinds import SCHEME

ADDED, MODIFIED, DELETED = "ADDED", "MODIFIED", "DELETED"
_log = logging.getLogger("kubedl.apiserver")


def _merge_patch(target, patch):
    """RFC 7386 JSON merge patch (returns a new tree, shares unpatched subtrees with `target`)."""
    if not isinstance(patch, dict):
        return patch
    if not isinstance(target, dict):
        target = {}
    out = dict(target)
    for k, v in patch.items():
        if v is None:
            out.pop(k, None)
        else:
            out[k] = _merge_patch(out.get(k), v)
    return out


class _Frozen:
    """Mix-in that turns every mutator of dict / list into an error (KUBEDL_FREEZE_SNAPSHOTS=1)."""

    def _ro(self, *a, **k):
        raise TypeError("a shared watch snapshot was modified by a handler registered with shared=True")
    __setitem__ = __delitem__ = clear = pop = popitem = setdefault = update = _ro
    append = extend = insert = remove = sort = reverse = __iadd__ = __imul__ = _ro


class _FrozenDict(_Frozen, dict):
    def __deepcopy__(self, memo):            # a copy is an ordinary, writable dict again
        return {k: copy.deepcopy(v, memo) for k, v in self.items()}


class _FrozenList(_Frozen, list):
    def __deepcopy__(self, memo):
        return [copy.deepcopy(v, memo) for v in self]


def _freeze(o):
    """Deep read-only view (a copy) of a JSON tree: what shared watchers get when KUBEDL_FREEZE_SNAPSHOTS=1, so that a
    handler that breaks the shared=True promise fails loudly in the tests instead of corrupting its neighbours."""
    if isinstance(o, dict):
        return _FrozenDict({k: _freeze(v) for k, v in o.items()})
    if isinstance(o, list):
        return _FrozenList([_freeze(v) for v in o])
    return o


_FREEZE_SNAPSHOTS = __import__("os").environ.get("KUBEDL_FREEZE_SNAPSHOTS", "0") == "1"


class ApiError(Exception):
    reason = "InternalError"
    code = 500


class NotFound(ApiError):
    reason, code = "NotFound", 404


class AlreadyExists(ApiError):
    reason, code = "AlreadyExists", 409


class Conflict(ApiError):
    reason, code = "Conflict", 409


class Invalid(ApiError):
    reason, code = "Invalid", 422


class Timeout(ApiError):
    reason, code = "Timeout", 504


class Gone(ApiError):
    """watch from a resourceVersion that has fallen out of the event history (HTTP 410)."""
    reason, code = "Expired", 410


def is_not_found(e: Exception) -> bool: return isinstance(e, NotFound)
def is_already_exists(e: Exception) -> bool: return isinstance(e, AlreadyExists)
def is_conflict(e: Exception) -> bool: return isinstance(e, Conflict)
def is_timeout(e: Exception) -> bool: return isinstance(e, Timeout)


WatchHandler = Callable[[str, dict, Optional[dict]], None]  # (event_type, obj, old_obj)


class APIServer:
    def __init__(self, scheme=SCHEME, install_all_crds: bool = True, history: int = 8192):
        self.scheme = scheme
        self._lock = threading.RLock()
        self._objs: Dict[str, Dict[Tuple[str, str], dict]] = collections.defaultdict(dict)
        self._rv = 0
        self._watchers: Dict[str, List[WatchHandler]] = collections.defaultdict(list)
        self._shared_handlers: set = set()        # handlers that take read-only snapshots (watch(shared=True))
        self._pending: Deque[Tuple[str, str, dict, Optional[dict]]] = collections.deque()
        self._dispatch_lock = threading.RLock()
        self._dispatching = threading.local()
        self._crds: Dict[str, dict] = {}
        # bounded event history so remote watchers can resume from a resourceVersion
        self._history: Deque[Tuple[int, str, str, dict]] = collections.deque(maxlen=history)
        self._history_floor = 0    # events with rv <= floor have been dropped
        self.fail_next: Dict[Tuple[str, str], Exception] = {}  # fault injection: (verb, kind) -> error
        self.request_log: Deque[Tuple[str, str, str]] = collections.deque(maxlen=65536)   # (verb, kind, ns/name), recent
        self.validate_schema = True      # OpenAPI validation of workload kinds on create / update
        self.event_ttl = 3600.0          # Events are dropped after an hour, as kube-apiserver does
        self._events_since_prune = 0
        if install_all_crds:
            for info in scheme.workload_kinds():
                self.install_crd(info.kind)

    # ------------------------------------------------------------------ discovery / CRDs
    def install_crd(self, kind: str) -> None:
        info = self.scheme.info(kind)
        self._crds[kind] = {"group": info.group, "version": info.version, "plural": info.plural}

    def uninstall_crd(self, kind: str) -> None:
        self._crds.pop(kind, None)

    def crd_installed(self, kind: str) -> bool:
        """Discovery: is the kind served? (reference workload_gate.go:90-106)"""
        return kind in self._crds

    def server_resources_for_group_version(self, gv: str) -> List[str]:
        return [k for k, v in self._crds.items() if f"{v['group']}/{v['version']}" == gv]

    # ------------------------------------------------------------------ helpers
    def _next_rv(self) -> str:
        self._rv += 1
        return str(self._rv)

    def current_rv(self) -> str:
        with self._lock:
            return str(self._rv)

    def _record(self, kind: str, ev: str, obj: dict) -> None:
        """Append to the watch history (caller holds the lock; obj is already a private copy)."""
        if self._history.maxlen is not None and len(self._history) == self._history.maxlen:
            self._history_floor = self._history[0][0]
        self._history.append((int(obj["metadata"]["resourceVersion"]), kind, ev, obj))

    def events_since(self, kind: str, rv: str) -> List[Tuple[str, dict]]:
        """Events of `kind` newer than resourceVersion `rv`, oldest first; `Gone` when the
        history no longer reaches back that far (the client must list again)."""
        since = int(rv or 0)
        with self._lock:
            if since < self._history_floor:
                raise Gone(f"too old resource version: {since} ({self._history_floor})")
            return [(ev, _copy(o)) for r, k, ev, o in self._history if k == kind and r > since]

    def _maybe_fail(self, verb: str, kind: str) -> None:
        err = self.fail_next.pop((verb, kind), None)
        if err is not None:
            raise err

    def _check_kind(self, kind: str) -> None:
        if self.scheme.is_workload(kind) and kind not in self._crds:
            raise NotFound(f"the server could not find the requested resource ({kind})")

    @staticmethod
    def _key(obj: dict) -> Tuple[str, str]:
        return (c.namespace_of(obj), c.name_of(obj))

    # ------------------------------------------------------------------ verbs
    def _validate(self, obj: dict) -> None:
        """What kube-apiserver does with the CRD's OpenAPI schema: reject a workload object whose
        shape is wrong (422 Invalid) instead of letting it reach a controller."""
        if not self.validate_schema or not self.scheme.is_workload(obj.get("kind", "")):
            return
        try:
            from ..api import crd as crd_mod
            crd_mod.validate(obj)
        except ImportError:          # jsonschema not installed: accept, as a server without a schema would
            return
        except Exception as e:
            msg = getattr(e, "message", None) or str(e)
            where = ".".join(str(p) for p in getattr(e, "absolute_path", []) or [])
            raise Invalid(f'{obj.get("kind")} "{c.name_of(obj)}" is invalid: {where + ": " if where else ""}{msg}')

    def create(self, obj: dict) -> dict:
        kind = obj.get("kind")
        if not kind or not c.name_of(obj):
            raise Invalid("object needs kind and metadata.name")
        self._check_kind(kind)
        self._validate(obj)
        with self._lock:
            self._maybe_fail("create", kind)
            key = self._key(obj)
            self.request_log.append(("create", kind, f"{key[0]}/{key[1]}"))
            if key in self._objs[kind]:
                raise AlreadyExists(f'{kind.lower()} "{key[1]}" already exists')
            o = _copy(obj)
            o.setdefault("apiVersion", self.scheme.api_version(kind) if self.scheme.knows(kind) else "v1")
            m = o.setdefault("metadata", {})
            m["namespace"] = key[0]
            m["uid"] = m.get("uid") or str(uuid.uuid4())
            m["resourceVersion"] = self._next_rv()
            m["creationTimestamp"] = m.get("creationTimestamp") or c.now_precise()
            m.setdefault("generation", 1)
            self._objs[kind][key] = o
            out = _copy(o)
            self._record(kind, ADDED, out)
            self._pending.append((kind, ADDED, out, None))
        self._dispatch()
        return _copy(out)

    def get(self, kind: str, namespace: str, name: str) -> dict:
        with self._lock:
            o = self._objs[kind].get((namespace or "default", name))
            if o is None:
                raise NotFound(f'{kind.lower()} "{name}" not found')
            return _copy(o)

    def list(self, kind: str, namespace: Optional[str] = None,
             label_selector: Optional[Dict[str, str]] = None) -> List[dict]:
        with self._lock:
            out = []
            for (ns, _), o in self._objs[kind].items():
                if namespace and ns != namespace:
                    continue
                if label_selector:
                    lab = o.get("metadata", {}).get("labels") or {}
                    if any(lab.get(k) != v for k, v in label_selector.items()):
                        continue
                out.append(_copy(o))
            out.sort(key=lambda o: (c.namespace_of(o), c.name_of(o)))
            return out

    def list_with_rv(self, kind: str, namespace: Optional[str] = None,
                     label_selector: Optional[Dict[str, str]] = None) -> Tuple[List[dict], str]:
        """LIST plus the resourceVersion it is consistent with (what a reflector watches from)."""
        with self._lock:
            return self.list(kind, namespace, label_selector), str(self._rv)

    def request_log_snapshot(self) -> List[Tuple[str, str, str]]:
        """A consistent copy of the recent (verb, kind, ns/name) request log; other threads keep appending."""
        with self._lock:
            return list(self.request_log)

    def _update(self, obj: dict, status_only: bool, spec_only: bool) -> dict:
        kind = obj.get("kind")
        with self._lock:
            self._maybe_fail("update_status" if status_only else "update", kind)
            key = self._key(obj)
            self.request_log.append(("update_status" if status_only else "update", kind, f"{key[0]}/{key[1]}"))
            cur = self._objs[kind].get(key)
            if cur is None:
                raise NotFound(f'{kind.lower()} "{key[1]}" not found')
            rv = obj.get("metadata", {}).get("resourceVersion")
            if rv and rv != cur["metadata"]["resourceVersion"]:
                raise Conflict(f"Operation cannot be fulfilled on {kind.lower()} \"{key[1]}\": the object "
                               f"has been modified; please apply your changes to the latest version")
            old = _copy(cur)
            new = _copy(cur)
            if status_only:
                new["status"] = _copy(obj.get("status", {}))
            else:
                incoming = _copy(obj)
                keep_meta = {k: cur["metadata"][k] for k in ("uid", "creationTimestamp", "namespace") if k in cur["metadata"]}
                if spec_only:
                    if "status" in cur:
                        incoming["status"] = _copy(cur["status"])
                    else:
                        incoming.pop("status", None)
                new = incoming
                new.setdefault("metadata", {}).update(keep_meta)
                if new.get("spec") != cur.get("spec"):
                    new["metadata"]["generation"] = int(cur["metadata"].get("generation", 1)) + 1
                else:
                    new["metadata"]["generation"] = cur["metadata"].get("generation", 1)
            new["metadata"]["resourceVersion"] = self._next_rv()
            self._objs[kind][key] = new
            out = _copy(new)
            self._record(kind, MODIFIED, out)
            self._pending.append((kind, MODIFIED, out, old))
        self._dispatch()
        return _copy(out)

    def update(self, obj: dict) -> dict:
        """PUT of the main resource; for workload kinds .status is ignored (status subresource)."""
        self._validate({k: v for k, v in obj.items() if k != "status"})
        return self._update(obj, status_only=False, spec_only=self.scheme.is_workload(obj.get("kind", "")))

    def update_status(self, obj: dict) -> dict:
        """PUT of the /status subresource: only .status changes."""
        return self._update(obj, status_only=True, spec_only=False)

    def patch_labels(self, kind: str, namespace: str, name: str, labels: Dict[str, Optional[str]]) -> dict:
        o = self.get(kind, namespace, name)
        lab = o.setdefault("metadata", {}).setdefault("labels", {})
        for k, v in labels.items():
            if v is None:
                lab.pop(k, None)
            else:
                lab[k] = v
        o["metadata"].pop("resourceVersion", None)
        return self._update(o, status_only=False, spec_only=False)

    def merge_patch(self, kind: str, namespace: str, name: str, patch: dict) -> dict:
        """RFC 7386 merge patch of the whole object, status included: what the REST façade's PATCH does, in one call
        instead of a get and an update, and atomic — the merge is applied to the version it was computed from
        (compare-and-swap on resourceVersion, retried), unless the patch itself pins a resourceVersion."""
        pinned = (patch.get("metadata") or {}).get("resourceVersion")
        for _ in range(16):
            with self._lock:
                cur = self._objs[kind].get((namespace or "default", name))
                if cur is None:
                    raise NotFound(f'{kind.lower()} "{name}" not found')
                new = _merge_patch(cur, patch)
                new["metadata"] = dict(new.get("metadata") or {})
                new["metadata"]["resourceVersion"] = pinned or cur["metadata"]["resourceVersion"]
            try:
                return self._update(new, status_only=False, spec_only=False)
            except Conflict:
                if pinned:
                    raise
        raise Conflict(f"{kind.lower()} \"{name}\" kept changing under a merge patch")

    def delete(self, kind: str, namespace: str, name: str, cascade: bool = True) -> None:
        with self._lock:
            self._maybe_fail("delete", kind)
            key = (namespace or "default", name)
            self.request_log.append(("delete", kind, f"{key[0]}/{key[1]}"))
            cur = self._objs[kind].pop(key, None)
            if cur is None:
                raise NotFound(f'{kind.lower()} "{name}" not found')
            gone = _copy(cur)
            gone.setdefault("metadata", {})["deletionTimestamp"] = c.now_rfc3339()
            gone["metadata"]["resourceVersion"] = self._next_rv()   # a deletion is a write too
            self._record(kind, DELETED, gone)
            self._pending.append((kind, DELETED, gone, None))
            dependents = []
            if cascade:
                uid = cur["metadata"].get("uid")
                for k2, objs in self._objs.items():
                    for key2, o2 in objs.items():
                        for ref in o2.get("metadata", {}).get("ownerReferences") or []:
                            if ref.get("uid") == uid:
                                dependents.append((k2, key2[0], key2[1]))
                                break
        self._dispatch()
        for k2, ns2, n2 in dependents:
            try:
                self.delete(k2, ns2, n2, cascade=True)  # background garbage collection
            except NotFound:
                pass

    # ------------------------------------------------------------------ watch
    def watch(self, kind: str, handler: WatchHandler, replay: bool = True, shared: bool = False) -> Callable[[], None]:
        """Register a watch.  With replay=True the handler first receives ADDED for every
        existing object (informer LIST+WATCH).  Returns an unsubscribe function.

        shared=True is a promise that the handler only READS (or stores, never mutates) the objects it is handed:
        all such handlers of an event get the same snapshot instead of a private copy each.  With six watchers on
        Pod (four controllers, the informer cache, the node agent) the private copies were a third of all object
        copies made for a job; client-go's informers make the same deal (objects from a shared informer are read-only)."""
        with self._lock:
            existing = [_copy(o) for o in self._objs[kind].values()] if replay else []
            self._watchers[kind].append(handler)
            if shared:
                self._shared_handlers.add(handler)
        for o in existing:
            handler(ADDED, o, None)

        def stop():
            with self._lock:
                if handler in self._watchers[kind]:
                    self._watchers[kind].remove(handler)
                self._shared_handlers.discard(handler)
        return stop

    def _dispatch(self) -> None:
        if getattr(self._dispatching, "active", False):
            return  # the outer frame of this thread drains the queue (keeps event order)
        with self._dispatch_lock:
            self._dispatching.active = True
            try:
                while True:
                    with self._lock:
                        if not self._pending:
                            break
                        kind, ev, obj, old = self._pending.popleft()
                        handlers = [(h, h in self._shared_handlers) for h in self._watchers.get(kind, ())]
                    snap = None
                    for h, shared in handlers:
                        try:
                            if shared:
                                if snap is None:         # one read-only snapshot for every handler that asked for it
                                    snap = (_copy(obj), _copy(old) if old is not None else None)
                                    if _FREEZE_SNAPSHOTS:
                                        snap = (_freeze(snap[0]), _freeze(snap[1]) if snap[1] is not None else None)
                                h(ev, snap[0], snap[1])
                            else:
                                h(ev, _copy(obj), _copy(old) if old is not None else None)
                        except Exception:
                            # a watcher's failure is its own: the write that caused the event has happened, and the
                            # writer (and the watchers after this one) must not pay for it
                            _log.error("watch handler for %s failed on %s %s/%s:\n%s", kind, ev, c.namespace_of(obj),
                                       c.name_of(obj), traceback.format_exc())
            finally:
                self._dispatching.active = False

    # ------------------------------------------------------------------ events
    def record_event(self, involved: dict, etype: str, reason: str, message: str, source: str = "kubedl") -> dict:
        ns = c.namespace_of(involved)
        name = f"{c.name_of(involved)}.{uuid.uuid4().hex[:12]}"
        now = c.now_rfc3339()
        ev = {"apiVersion": "v1", "kind": "Event", "metadata": {"name": name, "namespace": ns},
              "involvedObject": {"kind": involved.get("kind", ""), "namespace": ns, "name": c.name_of(involved),
                                 "uid": c.uid_of(involved), "apiVersion": involved.get("apiVersion", "")},
              "type": etype, "reason": reason, "message": message, "count": 1,
              "firstTimestamp": now, "lastTimestamp": now, "source": {"component": source}}
        out = self.create(ev)
        self._events_since_prune += 1
        if self._events_since_prune >= 256:
            self._events_since_prune = 0
            self.prune_events()
        return out

    def prune_events(self, now: Optional[float] = None) -> int:
        """Delete Events older than `event_ttl` (kube-apiserver's --event-ttl, 1 h by default)."""
        import time as _time
        cutoff = (now if now is not None else _time.time()) - self.event_ttl
        with self._lock:
            old = [(ns, n) for (ns, n), e in self._objs["Event"].items()
                   if (c.parse_time(e.get("lastTimestamp")) or 0.0) < cutoff]
        for ns, n in old:
            try:
                self.delete("Event", ns, n, cascade=False)
            except NotFound:
                pass
        return len(old)

    def events_for(self, involved: dict) -> List[dict]:
        uid = c.uid_of(involved)
        return [e for e in self.list("Event", c.namespace_of(involved))
                if e.get("involvedObject", {}).get("uid") == uid]

    # ------------------------------------------------------------------ convenience
    def apply(self, obj: dict) -> dict:
        """kubectl-apply-like: create, or replace spec/metadata if it exists."""
        try:
            return self.create(obj)
        except AlreadyExists:
            cur = self.get(obj["kind"], c.namespace_of(obj), c.name_of(obj))
            merged = _copy(obj)
            merged.setdefault("metadata", {}).pop("resourceVersion", None)
            if "status" in cur:
                merged["status"] = cur["status"]
            return self.update(merged)

    def kinds_present(self) -> List[str]:
        with self._lock:
            return [k for k, v in self._objs.items() if v]
