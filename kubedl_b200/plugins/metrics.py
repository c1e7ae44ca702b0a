"""Prometheus job metrics with the reference's names and labels (reference
pkg/metrics/job_metrics.go:33-194, status_counter.go:31-84, monitor.go:27-36, docs/metrics.md:7-16):

  kubedl_jobs_created / _deleted / _successful / _failed / _restarted {kind=<lower-case kind>}
  kubedl_jobs_running / kubedl_jobs_pending {kind}     computed on scrape by listing jobs
  kubedl_jobs_first_pod_launch_delay_seconds {kind,name,namespace,uid}   histogram
  kubedl_jobs_all_pods_launch_delay_seconds  {kind,name,namespace,uid}   histogram

Quirks kept: counters use the lower-cased kind, histograms the original-case kind
(job_metrics.go:76-77 vs :161-166); "pending" = the only condition is Created (:108-110).
Addition: raw (unbucketed) observations are also kept in `JobMetrics.launch_delays` so the
launch-delay benchmark can report p50/p90 (default buckets stop at 10 s).
"""
from __future__ import annotations

import collections
import os
import threading
from typing import Callable, Dict, List, Optional, Tuple

from prometheus_client import CollectorRegistry, Histogram, start_http_server
from prometheus_client.core import GaugeMetricFamily, Metric

from ..api import common as c
from ..utils import k8sutil
from ..utils import status as st


class _ExactCounter:
    def __init__(self, name: str, doc: str):
        self.name, self.doc = name, doc
        self._values: Dict[str, float] = {}
        self._lock = threading.Lock()

    def labels(self, kind: str) -> "_ExactCounterChild":
        with self._lock:
            self._values.setdefault(kind, 0.0)
        return _ExactCounterChild(self, kind)

    def metric(self) -> Metric:
        m = Metric(self.name, self.doc, "gauge")  # typed gauge so no _total suffix is appended
        with self._lock:
            for kind, v in sorted(self._values.items()):
                m.add_sample(self.name, {"kind": kind}, v)
        return m


class _ExactCounterChild:
    def __init__(self, parent: _ExactCounter, kind: str):
        self.parent, self.kind = parent, kind

    def inc(self, n: float = 1.0) -> None:
        with self.parent._lock:
            self.parent._values[self.kind] = self.parent._values.get(self.kind, 0.0) + n

    def value(self) -> float:
        with self.parent._lock:
            return self.parent._values.get(self.kind, 0.0)


class _Families:
    """One set of metric families per registry (the reference uses the process-wide default)."""

    def __init__(self, registry: CollectorRegistry):
        self.registry = registry
        # prometheus_client would rename a Counter to <name>_total; the Go client (and the
        # reference's dashboards, docs/metrics.md:7-11) use the bare names, so the counters are
        # kept here and exposed by collect() with the exact sample names.
        mk = lambda name, doc: _ExactCounter(name, doc)
        self.created = mk("kubedl_jobs_created", "Counts number of jobs created")
        self.deleted = mk("kubedl_jobs_deleted", "Counts number of jobs deleted")
        self.success = mk("kubedl_jobs_successful", "Counts number of jobs successfully finished")
        self.failure = mk("kubedl_jobs_failed", "Counts number of jobs failed")
        self.restart = mk("kubedl_jobs_restarted", "Counts number of jobs restarted")
        labels = ["kind", "name", "namespace", "uid"]
        self.first_pod = Histogram("kubedl_jobs_first_pod_launch_delay_seconds",
                                   "Histogram for recording launch delay duration(from job created to first pod running).",
                                   labels, registry=registry)
        self.all_pods = Histogram("kubedl_jobs_all_pods_launch_delay_seconds",
                                  "Histogram for recording sync launch delay duration(from job created to all pods running).",
                                  labels, registry=registry)
        self.gauge_sources: Dict[str, Callable[[], Tuple[int, int]]] = {}
        # The two histograms are labelled per JOB (name, namespace, uid), as in the reference: one pair of series per
        # job ever launched, never dropped there.  Here the oldest jobs' series go once more than
        # KUBEDL_LAUNCH_DELAY_SERIES_MAX jobs (2,000) hold one — ~14 KB per job for the two histograms' buckets
        # (scripts/soak_control_plane.py), and an operator runs for months.
        self.series_max = int(os.environ.get("KUBEDL_LAUNCH_DELAY_SERIES_MAX", "2000"))
        self._series: "collections.OrderedDict[Tuple[str, str, str, str], None]" = collections.OrderedDict()
        self._series_lock = threading.Lock()
        registry.register(self)

    def observe_delay(self, hist: Histogram, labels: Tuple[str, str, str, str], delay: float) -> None:
        hist.labels(*labels).observe(delay)
        with self._series_lock:
            self._series[labels] = None
            self._series.move_to_end(labels)
            drop = []
            while len(self._series) > self.series_max:
                drop.append(self._series.popitem(last=False)[0])
        for lv in drop:
            for h in (self.first_pod, self.all_pods):
                try:
                    h.remove(*lv)
                except KeyError:
                    pass

    # custom collector for the scrape-time gauges (promauto.NewGaugeFunc upstream)
    def collect(self):
        running = GaugeMetricFamily("kubedl_jobs_running", "Counts number of jobs running currently", labels=["kind"])
        pending = GaugeMetricFamily("kubedl_jobs_pending", "Counts number of jobs pending currently", labels=["kind"])
        for kind, fn in list(self.gauge_sources.items()):
            try:
                r, p = fn()
            except Exception:
                r, p = 0, 0
            running.add_metric([kind], r)
            pending.add_metric([kind], p)
        yield running
        yield pending
        for ctr in (self.created, self.deleted, self.success, self.failure, self.restart):
            yield ctr.metric()

    def describe(self):
        return []


_families: Dict[int, _Families] = {}
_lock = threading.Lock()
DEFAULT_REGISTRY = CollectorRegistry()


def _get_families(registry: CollectorRegistry) -> _Families:
    with _lock:
        f = _families.get(id(registry))
        if f is None:
            f = _Families(registry)
            _families[id(registry)] = f
        return f


def job_status_counter(kind: str, reader, predicate: Callable[[dict], bool]) -> int:
    """reference status_counter.go:31-48"""
    if hasattr(reader, "count"):          # the informer cache counts in place; predicates only read the status
        return reader.count(kind, lambda job: predicate(job.get("status") or {}))
    return sum(1 for job in reader.list(kind) if predicate(c.status_of(job)))


def _is_pending(status: dict) -> bool:
    return st.is_created(status) and len(status.get("conditions") or []) == 1


def _pod_ready_time(pod: dict) -> Optional[float]:
    for cond in (pod.get("status") or {}).get("conditions") or []:
        if cond.get("type") == "Ready":
            return c.parse_time(cond.get("lastTransitionTime"))
    return None


class _BoundedSet:
    """Insertion-ordered set that forgets its oldest members beyond `limit` (job UIDs seen)."""

    def __init__(self, limit: int):
        self.limit = limit
        self._d: "collections.OrderedDict" = collections.OrderedDict()

    def add(self, x) -> None:
        self._d[x] = None
        while len(self._d) > self.limit:
            self._d.popitem(last=False)

    def __contains__(self, x) -> bool:
        return x in self._d

    def __len__(self) -> int:
        return len(self._d)


class JobMetrics:
    def __init__(self, kind: str, reader, registry: Optional[CollectorRegistry] = None):
        self.kind = kind
        fam = _get_families(registry or DEFAULT_REGISTRY)
        self.registry = fam.registry
        lk = kind.lower()
        self._created, self._deleted = fam.created.labels(lk), fam.deleted.labels(lk)
        self._success, self._failure = fam.success.labels(lk), fam.failure.labels(lk)
        self._restart = fam.restart.labels(lk)
        self._first, self._all, self._fam = fam.first_pod, fam.all_pods, fam
        # bounded: an operator runs for months
        self.launch_delays: Dict[str, collections.deque] = {"first_pod": collections.deque(maxlen=100_000),
                                                            "all_pods": collections.deque(maxlen=100_000)}
        # The reference means to observe each delay once per job but its guards leak (first-pod is
        # re-observed on every pass while Created stays True, job.go:243-245; all-pods again after
        # a restart because Restarting has already flipped back to Running, job.go:255-259).
        # Observing once per job UID gives the intended semantics.
        self._seen_first = _BoundedSet(200_000)
        self._seen_all = _BoundedSet(200_000)
        if reader is not None:
            fam.gauge_sources[lk] = lambda: (job_status_counter(kind, reader, st.is_running),
                                             job_status_counter(kind, reader, _is_pending))

    def created_inc(self): self._created.inc()
    def deleted_inc(self): self._deleted.inc()
    def success_inc(self): self._success.inc()
    def failure_inc(self): self._failure.inc()
    def restart_inc(self): self._restart.inc()

    def first_pod_launch_delay_seconds(self, active_pods: List[dict], job: dict, status: dict) -> None:
        """min over Running pods of PodReady.lastTransitionTime − job creationTimestamp
        (reference job_metrics.go:139-167)."""
        if not st.is_running(status):
            return
        earliest = None
        for pod in active_pods:
            if k8sutil.phase_of(pod) != c.POD_RUNNING:
                continue
            t = _pod_ready_time(pod)
            if t is None:
                continue
            if earliest is None or t < earliest:
                earliest = t
        created = c.parse_time(job.get("metadata", {}).get("creationTimestamp"))
        if earliest is None or created is None:
            return
        if c.uid_of(job) in self._seen_first:
            return
        self._seen_first.add(c.uid_of(job))
        delay = earliest - created
        self._fam.observe_delay(self._first, (self.kind, c.name_of(job), c.namespace_of(job), c.uid_of(job)), delay)
        self.launch_delays["first_pod"].append((c.key_of(job), delay))

    def all_pods_launch_delay_seconds(self, pods: List[dict], job: dict, status: dict) -> None:
        """max PodReady.lastTransitionTime − creationTimestamp; aborts if any pod is not Running
        (reference job_metrics.go:169-194)."""
        if not st.is_running(status) or not status.get("startTime"):
            return
        created = c.parse_time(job.get("metadata", {}).get("creationTimestamp"))
        if created is None:
            return
        final = created
        for pod in pods:
            if k8sutil.phase_of(pod) != c.POD_RUNNING:
                return
            t = _pod_ready_time(pod)
            if t is not None and t > final:
                final = t
        if c.uid_of(job) in self._seen_all:
            return
        self._seen_all.add(c.uid_of(job))
        delay = final - created
        self._fam.observe_delay(self._all, (self.kind, c.name_of(job), c.namespace_of(job), c.uid_of(job)), delay)
        self.launch_delays["all_pods"].append((c.key_of(job), delay))


def start_monitoring(port: int, registry: Optional[CollectorRegistry] = None, addr: str = "0.0.0.0"):
    """HTTP /metrics for the kubedl_* registry on --metrics-addr (reference monitor.go:27-36)."""
    return start_http_server(port, addr=addr, registry=registry or DEFAULT_REGISTRY)


def counter_value(registry: CollectorRegistry, name: str, labels: Dict[str, str]) -> float:
    return registry.get_sample_value(name, labels) or 0.0
