"""controller-runtime equivalents: Client (cached reads), EventRecorder, Controller, Manager.

Mirrors what main.go builds (reference main.go:48-115): a manager holding the scheme, a cached
client, an event recorder factory, N controllers each with a rate-limited work queue and
`MaxConcurrentReconciles` workers, optional leader election, and start/stop.  Watches deliver
(event, obj, old) to per-controller handlers which map events to reconcile `Request`s —
`EnqueueRequestForOwner` and predicate functions are expressed as plain callables.

Two ways to run: `start()` spawns worker threads (production, `cmd.manager`); `run_until_idle()`
drains every queue on the calling thread (deterministic tests — the stronger harness SURVEY.md
§4 asks for, since the reference never runs a reconcile loop in its tests).
"""
from __future__ import annotations

import collections
import copy  # noqa: F401
import fcntl
import os
import queue
import threading
import time
import traceback
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Tuple

from ..api import common as c
from ..api.common import deepcopy as _copy
from ..api.kinds import SCHEME
from ..apiserver.store import ADDED, DELETED, MODIFIED, APIServer, NotFound
from ..utils.logger import logger
from .workqueue import RateLimitingQueue

log = logger("kubedl.runtime")


@dataclass(frozen=True)
class Request:
    namespace: str
    name: str

    @property
    def key(self) -> str:
        return f"{self.namespace}/{self.name}"


@dataclass
class Result:
    requeue: bool = False
    requeue_after: float = 0.0


def _older(a: Optional[dict], b: Optional[dict], unknown: bool = False) -> bool:
    """Is `a` an older version of the object than `b`?  resourceVersions are compared as the
    integers every real API server uses; if they are not integers the answer is `unknown`."""
    if a is None or b is None:
        return False
    try:
        return int(a["metadata"]["resourceVersion"]) < int(b["metadata"]["resourceVersion"])
    except (KeyError, TypeError, ValueError):
        return unknown


class Client:
    """Reads come from an informer cache kept current by watches; writes go to the API server.
    `reader()` gives uncached reads (reference pkg/util/runtime.go:29-48)."""

    def __init__(self, api: APIServer):
        self.api = api
        self._cache: Dict[str, Dict[Tuple[str, str], dict]] = {}
        # kind -> job-name label value -> keys of the cached objects carrying it.  Every reconcile lists the pods and
        # services of ONE job by {group-name, job-name}; without the index that is a walk over every pod of the
        # namespace (8 -> 22 ms per job going from 200 to 800 live jobs).
        self._by_job: Dict[str, Dict[str, set]] = {}
        self._lock = threading.RLock()
        self._stops: List[Callable[[], None]] = []

    def ensure_informer(self, kind: str) -> None:
        with self._lock:
            if kind in self._cache:
                return
            self._cache[kind] = {}
            self._by_job[kind] = {}

        def on_event(ev, obj, old):
            key = (c.namespace_of(obj), c.name_of(obj))
            with self._lock:
                if ev == DELETED:
                    self._store(kind, key, None)
                elif not _older(obj, self._cache[kind].get(key)):
                    self._store(kind, key, obj)
        self._stops.append(self.api.watch(kind, on_event, replay=True, shared=True))   # cache entries are never mutated

    def _store(self, kind: str, key: Tuple[str, str], obj: Optional[dict]) -> None:
        """Put (or, with None, drop) a cache entry and keep the job-name index in step.  Caller holds the lock."""
        cache, index = self._cache[kind], self._by_job[kind]
        cur = cache.get(key)
        if cur is not None:
            job = (cur.get("metadata", {}).get("labels") or {}).get(c.JOB_NAME_LABEL)
            if job is not None:
                keys = index.get(job)
                if keys is not None:
                    keys.discard(key)
                    if not keys:
                        del index[job]
        if obj is None:
            cache.pop(key, None)
            return
        cache[key] = obj
        job = (obj.get("metadata", {}).get("labels") or {}).get(c.JOB_NAME_LABEL)
        if job is not None:
            index.setdefault(job, set()).add(key)

    def _remember(self, obj: dict) -> dict:
        """Read-your-writes: what a write returned goes into the cache at once if it is newer than
        what the watch has delivered so far.  Against a remote API server the watch event for our
        own write arrives later; a reconcile in between would otherwise compute the next status
        from a copy that lacks the previous one (and, writing it, drop a condition)."""
        kind = obj.get("kind", "")
        with self._lock:
            cache = self._cache.get(kind)
            if cache is not None:
                key = (c.namespace_of(obj), c.name_of(obj))
                cur = cache.get(key)
                if cur is not None and _older(cur, obj):   # never resurrect what a DELETED event removed
                    self._store(kind, key, _copy(obj))
        return obj

    def get(self, kind: str, namespace: str, name: str) -> dict:
        self.ensure_informer(kind)
        with self._lock:
            o = self._cache[kind].get((namespace or "default", name))
            if o is None:
                raise NotFound(f'{kind.lower()} "{name}" not found')
            return _copy(o)

    def list(self, kind: str, namespace: Optional[str] = None,
             label_selector: Optional[Dict[str, str]] = None) -> List[dict]:
        self.ensure_informer(kind)
        with self._lock:
            out = []
            cache = self._cache[kind]
            job = (label_selector or {}).get(c.JOB_NAME_LABEL)
            if job is not None:      # one job's objects: through the index instead of over the whole cache
                items = [(k, cache[k]) for k in self._by_job[kind].get(job, ()) if k in cache]
            else:
                items = list(cache.items())
            for (ns, _), o in items:
                if namespace and ns != namespace:
                    continue
                if label_selector:
                    lab = o.get("metadata", {}).get("labels") or {}
                    if any(lab.get(k) != v for k, v in label_selector.items()):
                        continue
                out.append(_copy(o))
            out.sort(key=lambda o: (c.namespace_of(o), c.name_of(o)))
            return out

    def count(self, kind: str, predicate: Callable[[dict], bool]) -> int:
        """How many cached objects of `kind` satisfy a READ-ONLY predicate — no copies (list() copies every object,
        and the running / pending gauges ask about every job of every kind on each scrape)."""
        self.ensure_informer(kind)
        with self._lock:
            return sum(1 for o in self._cache[kind].values() if predicate(o))

    def reader(self) -> APIServer:
        return self.api

    # writes
    def create(self, obj: dict) -> dict: return self.api.create(obj)
    def update(self, obj: dict) -> dict: return self._remember(self.api.update(obj))
    def update_status(self, obj: dict) -> dict: return self._remember(self.api.update_status(obj))
    def delete(self, kind: str, namespace: str, name: str) -> None: self.api.delete(kind, namespace, name)

    def stop(self) -> None:
        for s in self._stops:
            s()
        self._stops.clear()


class EventRecorder:
    """record.EventRecorder: Events are stored in the API server and kept in memory."""

    def __init__(self, api: APIServer, component: str):
        self.api, self.component = api, component
        self.events: "collections.deque" = collections.deque(maxlen=4096)  # (object key, type, reason, message), most recent
        self._lock = threading.Lock()
        # Against a remote API server an Event is a round trip; client-go's recorder hands events
        # to a broadcaster goroutine so that reconciles never wait for them, and so do we.
        self._async = bool(getattr(api, "remote", False))
        self._q: "queue.Queue" = queue.Queue(maxsize=1000)
        self._thread: Optional[threading.Thread] = None

    def _drain(self) -> None:
        while True:
            item = self._q.get()
            if item is None:
                return
            try:
                self.api.record_event(*item, source=self.component)
            except Exception:  # events are best effort
                pass
            finally:
                self._q.task_done()

    def flush(self, timeout: float = 5.0) -> None:
        """Wait until queued events have been written (tests, shutdown)."""
        deadline = time.monotonic() + timeout
        while self._q.unfinished_tasks and time.monotonic() < deadline:
            time.sleep(0.005)

    def event(self, obj: dict, etype: str, reason: str, message: str) -> None:
        with self._lock:
            self.events.append((c.key_of(obj), etype, reason, message))
            if self._async and self._thread is None:
                self._thread = threading.Thread(target=self._drain, name=f"events-{self.component}", daemon=True)
                self._thread.start()
        if self._async:
            try:
                self._q.put_nowait((_copy({k: obj.get(k) for k in ("kind", "apiVersion", "metadata")}),
                                    etype, reason, message))
            except queue.Full:       # dropping beats blocking a reconcile (client-go does the same)
                pass
            return
        try:
            self.api.record_event(obj, etype, reason, message, source=self.component)
        except Exception:  # events are best effort
            pass

    def eventf(self, obj: dict, etype: str, reason: str, fmt: str, *args) -> None:
        self.event(obj, etype, reason, fmt % args if args else fmt)

    def reasons(self) -> List[str]:
        with self._lock:
            return [e[2] for e in self.events]


EventHandler = Callable[[str, dict, Optional[dict]], List[Request]]


def enqueue_request_for_object(ev: str, obj: dict, old: Optional[dict]) -> List[Request]:
    return [Request(c.namespace_of(obj), c.name_of(obj))]


def enqueue_request_for_owner(owner_kind: str, is_controller: bool = True) -> EventHandler:
    """handler.EnqueueRequestForOwner{OwnerType, IsController}."""
    def handler(ev: str, obj: dict, old: Optional[dict]) -> List[Request]:
        out = []
        for o in (obj, old):
            if o is None:
                continue
            for ref in o.get("metadata", {}).get("ownerReferences") or []:
                if ref.get("kind") != owner_kind:
                    continue
                if is_controller and not ref.get("controller"):
                    continue
                r = Request(c.namespace_of(o), ref.get("name", ""))
                if r not in out:
                    out.append(r)
        return out
    return handler


@dataclass
class Predicates:
    """predicate.Funcs: return False to drop the event before it reaches the handler."""
    create: Callable[[dict], bool] = lambda obj: True
    update: Callable[[dict, dict], bool] = lambda old, new: True
    delete: Callable[[dict], bool] = lambda obj: True


class Controller:
    def __init__(self, name: str, reconciler: Callable[[Request], Result], manager: "Manager",
                 max_concurrent_reconciles: int = 1):
        self.name = name
        self.reconciler = reconciler
        self.manager = manager
        self.max_concurrent = max(1, int(max_concurrent_reconciles))
        self.queue = RateLimitingQueue(name)
        self._watches: List[Tuple[str, EventHandler, Predicates]] = []
        self._stops: List[Callable[[], None]] = []
        self._threads: List[threading.Thread] = []
        self.reconcile_count = 0
        self.error_count = 0
        self.trace: List[dict] = []  # per-reconcile span log

    def watch(self, kind: str, handler: EventHandler = enqueue_request_for_object,
              predicates: Optional[Predicates] = None) -> None:
        self._watches.append((kind, handler, predicates or Predicates()))

    def _on_event(self, handler: EventHandler, pred: Predicates):
        def cb(ev: str, obj: dict, old: Optional[dict]):
            try:
                if ev == ADDED and not pred.create(obj):
                    return
                if ev == MODIFIED and not pred.update(old if old is not None else obj, obj):
                    return
                if ev == DELETED and not pred.delete(obj):
                    return
                for req in handler(ev, obj, old):
                    if req.name:
                        self.queue.add(req)
            except Exception:
                log.error("event handler of %s failed: %s", self.name, traceback.format_exc())
        return cb

    def start_watches(self) -> None:
        for kind, handler, pred in self._watches:
            self.manager.client.ensure_informer(kind)
            # predicates and request mappers only read the objects of an event
            self._stops.append(self.manager.api.watch(kind, self._on_event(handler, pred), replay=True, shared=True))

    def process_one(self, block: bool = False, timeout: Optional[float] = None) -> bool:
        req, shutdown = self.queue.get(block=block, timeout=timeout)
        if req is None:
            return False
        t0 = time.perf_counter()
        err = None
        res = Result()
        try:
            res = self.reconciler(req) or Result()
        except Exception as e:  # reconcile error -> rate-limited requeue
            err = e
            from ..apiserver.store import Conflict
            if isinstance(e, Conflict):
                # optimistic concurrency doing its job (a write computed from a stale read was refused):
                # routine, requeued like any error but not reported as one
                self.conflict_count = getattr(self, "conflict_count", 0) + 1
                log.debug("reconcile %s %s: %s; requeued", self.name, req.key, e)
            else:
                self.error_count += 1
                log.warning("reconcile %s %s failed: %s", self.name, req.key, e)
                log.debug(traceback.format_exc())
        finally:
            self.reconcile_count += 1
            self.trace.append({"controller": self.name, "key": req.key, "seconds": time.perf_counter() - t0,
                               "error": repr(err) if err else None})
            if len(self.trace) > 4096:
                del self.trace[:2048]
        cm = self.manager.controller_metrics
        if cm is not None:
            result = "error" if err is not None else "requeue_after" if res.requeue_after > 0 else \
                "requeue" if res.requeue else "success"
            cm.observe_reconcile(self.name, result, time.perf_counter() - t0)
        if err is not None:
            self.queue.add_rate_limited(req)
        elif res.requeue_after > 0:
            self.queue.forget(req)
            self.queue.add_after(req, res.requeue_after)
        elif res.requeue:
            self.queue.add_rate_limited(req)
        else:
            self.queue.forget(req)
        self.queue.done(req)
        return True

    def _worker(self) -> None:
        while not self.manager.stopping.is_set():
            self.process_one(block=True, timeout=0.2)

    def start_workers(self) -> None:
        for i in range(self.max_concurrent):
            t = threading.Thread(target=self._worker, name=f"{self.name}-{i}", daemon=True)
            t.start()
            self._threads.append(t)

    def stop(self) -> None:
        for s in self._stops:
            s()
        self._stops.clear()
        self.queue.shutdown()


class LeaderElector:
    """Lease-based single-active-manager guard (reference main.go:56-57,73).  On one box the lease
    is an flock on a file; `acquire` blocks (or returns False when non-blocking)."""

    def __init__(self, lock_path: str = "/tmp/kubedl-b200-leader.lock", identity: Optional[str] = None):
        self.lock_path = lock_path
        self.identity = identity or f"{os.uname().nodename}_{os.getpid()}"
        self._fd = None

    def acquire(self, blocking: bool = True) -> bool:
        self._fd = os.open(self.lock_path, os.O_CREAT | os.O_RDWR, 0o644)
        try:
            fcntl.flock(self._fd, fcntl.LOCK_EX if blocking else fcntl.LOCK_EX | fcntl.LOCK_NB)
        except BlockingIOError:
            os.close(self._fd)
            self._fd = None
            return False
        os.ftruncate(self._fd, 0)
        os.write(self._fd, self.identity.encode())
        return True

    def release(self) -> None:
        if self._fd is not None:
            fcntl.flock(self._fd, fcntl.LOCK_UN)
            os.close(self._fd)
            self._fd = None


class LeaseLeaderElector:
    """Leader election on a `coordination.k8s.io/v1` Lease — what controller-runtime does for the
    reference when `--enable-leader-election` is set (reference main.go:56-57,73) — used when the
    API server is remote, where an flock on one machine means nothing.

    A candidate owns the lease if it created it, if the holder field is empty, or if the holder
    has not renewed within `leaseDurationSeconds`.  Every write carries the resourceVersion that
    was read, so two candidates racing for an expired lease cannot both win.  The holder renews
    every `retry_period`; if it cannot renew for `renew_deadline` it calls `on_lost` (the manager
    stops: two active operators would fight over the same pods)."""

    def __init__(self, api, namespace: str = "kubedl-system", name: str = "kubedl-election",
                 identity: Optional[str] = None, lease_duration: float = 15.0, renew_deadline: float = 10.0,
                 retry_period: float = 2.0, on_lost: Optional[Callable[[], None]] = None):
        self.api, self.namespace, self.name = api, namespace, name
        self.identity = identity or f"{os.uname().nodename}_{os.getpid()}_{os.urandom(3).hex()}"
        self.lease_duration, self.renew_deadline, self.retry_period = lease_duration, renew_deadline, retry_period
        self.on_lost = on_lost
        self.is_leader = False
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None

    def _spec(self, prev: Optional[dict]) -> dict:
        now = c.now_precise()
        prev = prev or {}
        mine = prev.get("holderIdentity") == self.identity
        return {"holderIdentity": self.identity, "leaseDurationSeconds": int(round(self.lease_duration)),
                "acquireTime": prev.get("acquireTime") if mine else now, "renewTime": now,
                "leaseTransitions": int(prev.get("leaseTransitions", 0)) + (0 if mine or not prev else 1)}

    def try_acquire_or_renew(self) -> bool:
        from ..apiserver.store import AlreadyExists, Conflict
        try:
            lease = self.api.get("Lease", self.namespace, self.name)
        except NotFound:
            try:
                self.api.create({"apiVersion": "coordination.k8s.io/v1", "kind": "Lease",
                                 "metadata": {"name": self.name, "namespace": self.namespace},
                                 "spec": self._spec(None)})
                return True
            except AlreadyExists:
                return False
        spec = lease.get("spec") or {}
        holder = spec.get("holderIdentity") or ""
        if holder and holder != self.identity:
            renewed = c.parse_time(spec.get("renewTime")) or 0.0
            if time.time() < renewed + float(spec.get("leaseDurationSeconds") or self.lease_duration):
                return False                     # somebody else holds a live lease
        lease["spec"] = self._spec(spec)
        try:
            self.api.update(lease)               # carries the resourceVersion we read
            return True
        except (Conflict, NotFound):
            return False

    def acquire(self, blocking: bool = True) -> bool:
        while not self._stop.is_set():
            try:
                if self.try_acquire_or_renew():
                    self.is_leader = True
                    self._thread = threading.Thread(target=self._renew_loop, name="leader-renew", daemon=True)
                    self._thread.start()
                    return True
            except Exception as e:
                log.warning("leader election: %r", e)
            if not blocking:
                return False
            self._stop.wait(self.retry_period)
        return False

    def _renew_loop(self) -> None:
        last_ok = time.monotonic()
        while not self._stop.wait(self.retry_period):
            ok = False
            try:
                ok = self.try_acquire_or_renew()
            except Exception as e:
                log.warning("leader election renew: %r", e)
            if ok:
                last_ok = time.monotonic()
            elif time.monotonic() - last_ok > self.renew_deadline:
                self.is_leader = False
                log.error("leader election lost (%s)", self.identity)
                if self.on_lost:
                    self.on_lost()
                return

    def release(self) -> None:
        self._stop.set()
        if self._thread is not None and self._thread is not threading.current_thread():
            self._thread.join(timeout=2)
        if not self.is_leader:
            return
        self.is_leader = False
        try:                                     # hand over at once instead of after leaseDuration
            lease = self.api.get("Lease", self.namespace, self.name)
            if (lease.get("spec") or {}).get("holderIdentity") == self.identity:
                lease["spec"]["holderIdentity"] = ""
                self.api.update(lease)
        except Exception:
            pass


class Manager:
    def __init__(self, api: Optional[APIServer] = None, scheme=SCHEME, leader_election: bool = False,
                 leader_lock_path: str = "/tmp/kubedl-b200-leader.lock",
                 leader_election_namespace: str = "kubedl-system"):
        self.api = api or APIServer(scheme)
        self.scheme = scheme
        self.client = Client(self.api)
        self.controllers: List[Controller] = []
        self.stopping = threading.Event()
        self._recorders: Dict[str, EventRecorder] = {}
        self.leader_election = leader_election
        if not leader_election:
            self.elector = None
        elif getattr(self.api, "remote", False):
            self.elector = LeaseLeaderElector(self.api, leader_election_namespace, on_lost=self.stop)
        else:
            self.elector = LeaderElector(leader_lock_path)
        self.runnables: List[Callable[["Manager"], None]] = []   # started with the manager (gpulet, metrics...)
        self.controller_metrics = None       # plugins/controller_metrics.ControllerMetrics (set_controller_metrics)
        self._started = False

    def set_controller_metrics(self, cm) -> None:
        """controller-runtime's registry: reconcile + workqueue + REST client instruments."""
        self.controller_metrics = cm
        for ctl in self.controllers:
            ctl.queue.metrics = cm.for_queue(ctl.name)
        if getattr(self.api, "remote", False):
            self.api.metrics = cm

    def get_event_recorder_for(self, name: str) -> EventRecorder:
        if name not in self._recorders:
            self._recorders[name] = EventRecorder(self.api, name)
        return self._recorders[name]

    def new_controller(self, name: str, reconciler: Callable[[Request], Result],
                       max_concurrent_reconciles: int = 1) -> Controller:
        ctl = Controller(name, reconciler, self, max_concurrent_reconciles)
        if self.controller_metrics is not None:
            ctl.queue.metrics = self.controller_metrics.for_queue(name)
        self.controllers.append(ctl)
        if self._started:
            ctl.start_watches()
        return ctl

    def add_runnable(self, fn: Callable[["Manager"], None]) -> None:
        self.runnables.append(fn)

    # ---- threaded mode
    def start(self, block: bool = False) -> None:
        if self.elector is not None:
            self.elector.acquire(blocking=True)
        self._started = True
        for ctl in self.controllers:
            ctl.start_watches()
        for ctl in self.controllers:
            ctl.start_workers()
        for fn in self.runnables:
            fn(self)
        if block:
            try:
                while not self.stopping.is_set():
                    time.sleep(0.2)
            except KeyboardInterrupt:
                pass
            self.stop()

    def stop(self) -> None:
        self.stopping.set()
        for ctl in self.controllers:
            ctl.stop()
        self.client.stop()
        if self.elector is not None:
            self.elector.release()
        if getattr(self.api, "remote", False):
            self.api.close()                     # ends the reflector threads

    # ---- deterministic mode (tests, single-threaded drivers)
    def start_watches(self) -> None:
        if not self._started:
            self._started = True
            for ctl in self.controllers:
                ctl.start_watches()

    def run_until_idle(self, max_iterations: int = 10000, advance_delayed: bool = False) -> int:
        """Process queued requests on this thread until every queue is empty.  Delayed items
        (add_after / rate-limited requeues) are waited for only when advance_delayed=True and
        their delay is short (< 50 ms)."""
        self.start_watches()
        n = 0
        for _ in range(max_iterations):
            progressed = False
            for ctl in self.controllers:
                while ctl.process_one(block=False):
                    progressed = True
                    n += 1
            if not progressed:
                if advance_delayed:
                    delays = [d for d in (ctl.queue.next_delay() for ctl in self.controllers) if d is not None]
                    if delays and min(delays) < 0.05:
                        time.sleep(max(min(delays), 0) + 0.001)
                        continue
                break
        return n
