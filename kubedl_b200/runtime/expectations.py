"""ControllerExpectations (k8s.io/kubernetes/pkg/controller semantics) — the guard against
acting on a stale cache: after issuing N creates the controller "expects" N ADDED events and
does not re-diff until they are observed or the expectation expires (5 min upstream).

The reference keys them `<jobKey>/<rtype>/pods` and `<jobKey>/<rtype>/services`
(pkg/job_controller/util.go:51-57) and ANDs them in SatisfyExpectations
(pkg/job_controller/expectations.go:11-27).  Note `expect_creations(key, n)` OVERWRITES the
record (SetExpectations), which the reference calls with n=1 per pod (pod.go:329-333): after
eight creates it expects one event, so the first ADDED that arrives lets a reconcile through
while seven pods are still unknown to the cache, and that reconcile creates them again
(AlreadyExists, error, rate-limited requeue).  Against an in-process store that never shows;
against a remote API server it showed on every other job.  The engine therefore counts with
`expect_one_more_creation`, which accumulates.
"""
from __future__ import annotations

import threading
import time
from typing import Callable, Dict, Optional

EXPECTATIONS_TIMEOUT = 5 * 60.0


class _Record:
    __slots__ = ("add", "dele", "ts")

    def __init__(self, add: int, dele: int, ts: float):
        self.add, self.dele, self.ts = add, dele, ts


class ControllerExpectations:
    def __init__(self, clock: Callable[[], float] = time.monotonic, ttl: float = EXPECTATIONS_TIMEOUT):
        self._lock = threading.Lock()
        self._store: Dict[str, _Record] = {}
        self.clock, self.ttl = clock, ttl

    def get(self, key: str) -> Optional[_Record]:
        with self._lock:
            return self._store.get(key)

    def set_expectations(self, key: str, add: int, dele: int) -> None:
        with self._lock:
            self._store[key] = _Record(add, dele, self.clock())

    def expect_creations(self, key: str, adds: int) -> None:
        self.set_expectations(key, adds, 0)

    def expect_deletions(self, key: str, dels: int) -> None:
        self.set_expectations(key, 0, dels)

    def expect_one_more_creation(self, key: str) -> None:
        """Count one more create in flight.  A live record with creates still unobserved grows;
        anything else (no record, satisfied, expired) starts again at one."""
        with self._lock:
            r = self._store.get(key)
            now = self.clock()
            if r is not None and r.add > 0 and now - r.ts <= self.ttl:
                r.add += 1
                r.ts = now
            else:
                self._store[key] = _Record(1, 0, now)

    def raise_expectations(self, key: str, add: int, dele: int) -> None:
        with self._lock:
            r = self._store.get(key)
            if r is not None:
                r.add += add
                r.dele += dele

    def lower_expectations(self, key: str, add: int, dele: int) -> None:
        with self._lock:
            r = self._store.get(key)
            if r is not None:
                r.add -= add
                r.dele -= dele

    def creation_observed(self, key: str) -> None:
        self.lower_expectations(key, 1, 0)

    def deletion_observed(self, key: str) -> None:
        self.lower_expectations(key, 0, 1)

    def satisfied_expectations(self, key: str) -> bool:
        with self._lock:
            r = self._store.get(key)
            if r is None:
                return True  # new controller / never set: sync
            if r.add <= 0 and r.dele <= 0:
                return True
            if self.clock() - r.ts > self.ttl:
                return True  # expired: something went wrong, resync
            return False

    def delete_expectations(self, key: str) -> None:
        with self._lock:
            self._store.pop(key, None)


def gen_expectation_pods_key(job_key: str, rtype: str) -> str:
    return f"{job_key}/{rtype.lower()}/pods"


def gen_expectation_services_key(job_key: str, rtype: str) -> str:
    return f"{job_key}/{rtype.lower()}/services"
