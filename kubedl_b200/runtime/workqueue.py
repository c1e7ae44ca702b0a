"""Rate-limited de-duplicating work queue (client-go workqueue semantics).

The reference uses `workqueue.NewNamedRateLimitingQueue(DefaultControllerRateLimiter())` twice:
as the reconcile queue inside controller-runtime and as `BackoffStatesQueue`, which only
counts retries per job (reference pkg/job_controller/job_controller.go:85-88, job.go:78-88,119).
Semantics kept: an item is never processed concurrently with itself (dirty / processing sets),
per-item exponential back-off 5 ms .. 1000 s, `forget` resets the failure count,
`num_requeues` reads it, `add_after` delays.
"""
from __future__ import annotations

import heapq
import threading
import time
from typing import Callable, Dict, Hashable, List, Optional, Set, Tuple


class ItemExponentialFailureRateLimiter:
    def __init__(self, base_delay: float = 0.005, max_delay: float = 1000.0):
        self.base, self.max = base_delay, max_delay
        self.failures: Dict[Hashable, int] = {}
        self._lock = threading.Lock()

    def when(self, item) -> float:
        with self._lock:
            n = self.failures.get(item, 0)
            self.failures[item] = n + 1
        return min(self.base * (2 ** n), self.max)

    def forget(self, item) -> None:
        with self._lock:
            self.failures.pop(item, None)

    def num_requeues(self, item) -> int:
        with self._lock:
            return self.failures.get(item, 0)


class RateLimitingQueue:
    def __init__(self, name: str = "", clock: Callable[[], float] = time.monotonic,
                 rate_limiter: Optional[ItemExponentialFailureRateLimiter] = None):
        self.name = name
        self.clock = clock
        self.limiter = rate_limiter or ItemExponentialFailureRateLimiter()
        self._cv = threading.Condition()
        self._queue: List[Hashable] = []
        self._dirty: Set[Hashable] = set()
        self._processing: Set[Hashable] = set()
        self._waiting: List[Tuple[float, int, Hashable]] = []  # (ready_at, seq, item)
        self._seq = 0
        self._shutdown = False
        self.adds = 0
        self.metrics = None      # plugins/controller_metrics.QueueMetrics, attached by the manager

    # ---- plain queue
    def add(self, item) -> None:
        with self._cv:
            if self._shutdown or item in self._dirty:
                return
            self.adds += 1
            self._dirty.add(item)
            if self.metrics is not None:
                self.metrics.add(item)
            if item in self._processing:
                return  # re-queued when done() is called
            self._queue.append(item)
            self._cv.notify()

    def _promote_ready(self) -> Optional[float]:
        """Move delayed items whose time has come; returns seconds until the next one."""
        now = self.clock()
        while self._waiting and self._waiting[0][0] <= now:
            _, _, item = heapq.heappop(self._waiting)
            if item not in self._dirty:
                self._dirty.add(item)
                if self.metrics is not None:
                    self.metrics.add(item)
                if item not in self._processing:
                    self._queue.append(item)
        return (self._waiting[0][0] - now) if self._waiting else None

    def get(self, block: bool = True, timeout: Optional[float] = None):
        """Returns (item, shutdown)."""
        deadline = None if timeout is None else self.clock() + timeout
        with self._cv:
            while True:
                wait_for = self._promote_ready()
                if self._queue:
                    item = self._queue.pop(0)
                    self._processing.add(item)
                    self._dirty.discard(item)
                    if self.metrics is not None:
                        self.metrics.get(item)
                    return item, False
                if self._shutdown:
                    return None, True
                if not block:
                    return None, False
                if deadline is not None:
                    rem = deadline - self.clock()
                    if rem <= 0:
                        return None, False
                    wait_for = rem if wait_for is None else min(wait_for, rem)
                self._cv.wait(wait_for if wait_for is None or wait_for > 0 else 0.001)

    def done(self, item) -> None:
        with self._cv:
            self._processing.discard(item)
            if self.metrics is not None:
                self.metrics.done(item)
            if item in self._dirty:
                self._queue.append(item)
                self._cv.notify()

    def __len__(self) -> int:
        with self._cv:
            return len(self._queue)

    def pending_delayed(self) -> int:
        with self._cv:
            return len(self._waiting)

    def next_delay(self) -> Optional[float]:
        with self._cv:
            return (self._waiting[0][0] - self.clock()) if self._waiting else None

    def idle(self) -> bool:
        with self._cv:
            self._promote_ready()
            return not self._queue and not self._processing

    # ---- delaying / rate limiting
    def add_after(self, item, delay: float) -> None:
        if delay <= 0:
            self.add(item)
            return
        with self._cv:
            if self._shutdown:
                return
            self._seq += 1
            heapq.heappush(self._waiting, (self.clock() + delay, self._seq, item))
            self._cv.notify()

    def add_rate_limited(self, item) -> None:
        if self.metrics is not None:
            self.metrics.retry()
        self.add_after(item, self.limiter.when(item))

    def forget(self, item) -> None:
        self.limiter.forget(item)

    def num_requeues(self, item) -> int:
        return self.limiter.num_requeues(item)

    def shutdown(self) -> None:
        with self._cv:
            self._shutdown = True
            self._cv.notify_all()
