"""The metrics controller-runtime and client-go publish for every operator built on them.

The reference serves two endpoints: the `kubedl_*` job metrics on
`--metrics-addr`, and — through `MetricsBindAddress` (reference main.go:70-72) — controller-runtime's
own registry on `--controller-metrics-addr` (`:8080`): reconcile counts / errors / latency per
controller, the client-go work-queue family per queue, and REST client request counts.  Dashboards
and alerts written for any kubebuilder operator query those names, so they are reproduced here
with the names and labels of controller-runtime v0.4 / client-go 0.16 (reference go.mod:53,63):

  controller_runtime_reconcile_total{controller,result}      result: success|error|requeue|requeue_after
  controller_runtime_reconcile_errors_total{controller}
  controller_runtime_reconcile_time_seconds{controller}      histogram
  workqueue_depth{name}  workqueue_adds_total{name}  workqueue_retries_total{name}
  workqueue_queue_duration_seconds{name}  workqueue_work_duration_seconds{name}      histograms
  workqueue_unfinished_work_seconds{name}  workqueue_longest_running_processor_seconds{name}
  rest_client_requests_total{code,method,host}  rest_client_request_latency_seconds{verb,url}
"""
from __future__ import annotations

import threading
import time
from http.server import ThreadingHTTPServer
from typing import Dict, Hashable, Optional, Tuple

from prometheus_client import CollectorRegistry, Counter, Gauge, Histogram
from prometheus_client.exposition import MetricsHandler

# client-go's exponential buckets for queue latencies: 1e-8 * 10^k, k = 0..9 (10 ns .. 10 s)
_QUEUE_BUCKETS = tuple(1e-8 * 10 ** k for k in range(10))
_RECONCILE_BUCKETS = (0.005, 0.01, 0.025, 0.05, 0.1, 0.25, 0.5, 1.0, 2.5, 5.0, 10.0)
_REST_BUCKETS = tuple(0.001 * 2 ** k for k in range(10))


class QueueMetrics:
    """Per-queue instrument bundle, driven by `RateLimitingQueue` (add / get / done / retry)."""

    def __init__(self, owner: "ControllerMetrics", name: str):
        self.name = name
        self._depth = owner.depth.labels(name)
        self._adds = owner.adds.labels(name)
        self._retries = owner.retries.labels(name)
        self._latency = owner.queue_duration.labels(name)
        self._work = owner.work_duration.labels(name)
        self._lock = threading.Lock()
        self._added_at: Dict[Hashable, float] = {}
        self._started_at: Dict[Hashable, float] = {}
        owner.unfinished.labels(name).set_function(self._unfinished)
        owner.longest.labels(name).set_function(self._longest)

    def add(self, item) -> None:
        self._adds.inc()
        self._depth.inc()
        with self._lock:
            self._added_at.setdefault(item, time.monotonic())

    def get(self, item) -> None:
        now = time.monotonic()
        self._depth.dec()
        with self._lock:
            t = self._added_at.pop(item, None)
            self._started_at[item] = now
        if t is not None:
            self._latency.observe(now - t)

    def done(self, item) -> None:
        with self._lock:
            t = self._started_at.pop(item, None)
        if t is not None:
            self._work.observe(time.monotonic() - t)

    def retry(self) -> None:
        self._retries.inc()

    def _unfinished(self) -> float:
        now = time.monotonic()
        with self._lock:
            return sum(now - t for t in self._started_at.values())

    def _longest(self) -> float:
        now = time.monotonic()
        with self._lock:
            return max((now - t for t in self._started_at.values()), default=0.0)


class ControllerMetrics:
    def __init__(self, registry: Optional[CollectorRegistry] = None):
        self.registry = registry if registry is not None else CollectorRegistry()
        r = self.registry
        self.reconcile_total = Counter("controller_runtime_reconcile_total", "Total number of reconciliations per controller",
                                       ["controller", "result"], registry=r)
        self.reconcile_errors = Counter("controller_runtime_reconcile_errors_total",
                                        "Total number of reconciliation errors per controller", ["controller"], registry=r)
        self.reconcile_time = Histogram("controller_runtime_reconcile_time_seconds", "Length of time per reconciliation per controller",
                                        ["controller"], buckets=_RECONCILE_BUCKETS, registry=r)
        self.depth = Gauge("workqueue_depth", "Current depth of workqueue", ["name"], registry=r)
        self.adds = Counter("workqueue_adds_total", "Total number of adds handled by workqueue", ["name"], registry=r)
        self.retries = Counter("workqueue_retries_total", "Total number of retries handled by workqueue", ["name"], registry=r)
        self.queue_duration = Histogram("workqueue_queue_duration_seconds",
                                        "How long in seconds an item stays in workqueue before being requested",
                                        ["name"], buckets=_QUEUE_BUCKETS, registry=r)
        self.work_duration = Histogram("workqueue_work_duration_seconds",
                                       "How long in seconds processing an item from workqueue takes.",
                                       ["name"], buckets=_QUEUE_BUCKETS, registry=r)
        self.unfinished = Gauge("workqueue_unfinished_work_seconds",
                                "How many seconds of work has been done that is in progress and hasn't been observed by work_duration.",
                                ["name"], registry=r)
        self.longest = Gauge("workqueue_longest_running_processor_seconds",
                             "How many seconds has the longest running processor for workqueue been running.", ["name"], registry=r)
        self.rest_requests = Counter("rest_client_requests_total", "Number of HTTP requests, partitioned by status code, method, and host.",
                                     ["code", "method", "host"], registry=r)
        self.rest_latency = Histogram("rest_client_request_latency_seconds", "Request latency in seconds. Broken down by verb and URL.",
                                      ["verb", "url"], buckets=_REST_BUCKETS, registry=r)
        self._queues: Dict[str, QueueMetrics] = {}

    # ------------------------------------------------------------------ hooks
    def for_queue(self, name: str) -> QueueMetrics:
        if name not in self._queues:
            self._queues[name] = QueueMetrics(self, name)
        return self._queues[name]

    def observe_reconcile(self, controller: str, result: str, seconds: float) -> None:
        self.reconcile_total.labels(controller, result).inc()
        if result == "error":
            self.reconcile_errors.labels(controller).inc()
        self.reconcile_time.labels(controller).observe(seconds)

    def observe_request(self, method: str, host: str, url_template: str, code: int, seconds: float) -> None:
        self.rest_requests.labels(str(code), method, host).inc()
        self.rest_latency.labels(method, url_template).observe(seconds)


def parse_bind_address(addr: str) -> Optional[Tuple[str, int]]:
    """`:8080` -> ("0.0.0.0", 8080); `127.0.0.1:9090`; `0` / `` -> None (disabled, as controller-runtime's "0")."""
    addr = (addr or "").strip()
    if addr in ("", "0"):
        return None
    host, _, port = addr.rpartition(":")
    return (host or "0.0.0.0", int(port))


def serve(metrics: ControllerMetrics, addr: str, manager=None) -> Optional[ThreadingHTTPServer]:
    """`/metrics`, plus — the reference has no tracing at all — `/debug/trace`: the last
    reconcile spans per controller (key, seconds, error) and `/debug/queues`: depth / delayed / busy per queue."""
    bind = parse_bind_address(addr)
    if bind is None:
        return None
    import json

    class Quiet(MetricsHandler.factory(metrics.registry)):
        def log_message(self, *a):
            pass

        def _json(self, obj) -> None:
            data = json.dumps(obj).encode()
            self.send_response(200)
            self.send_header("Content-Type", "application/json")
            self.send_header("Content-Length", str(len(data)))
            self.end_headers()
            self.wfile.write(data)

        def do_GET(self):
            path = self.path.split("?")[0]
            if manager is not None and path == "/debug/trace":
                return self._json({ctl.name: list(ctl.trace[-200:]) for ctl in manager.controllers})
            if manager is not None and path == "/debug/queues":
                return self._json({ctl.name: {"depth": len(ctl.queue), "delayed": ctl.queue.pending_delayed(),
                                              "idle": ctl.queue.idle(), "reconciles": ctl.reconcile_count,
                                              "errors": ctl.error_count} for ctl in manager.controllers})
            return super().do_GET()

    class Server(ThreadingHTTPServer):
        daemon_threads = True
        allow_reuse_address = True

        def shutdown(self):
            super().shutdown()
            self.server_close()

    srv = Server(bind, Quiet)
    threading.Thread(target=srv.serve_forever, name="controller-metrics", daemon=True).start()
    return srv
