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
        self._sweep_at = 1024

    def get(self, key: str) -> Optional[_Record]:
        with self._lock:
            return self._store.get(key)

    def set_expectations(self, key: str, add: int, dele: int) -> None:
        with self._lock:
            self._store[key] = _Record(add, dele, self.clock())
            grown = len(self._store) > self._sweep_at
        if grown:                                   # amortised: a sweep each time the table has doubled
            self.sweep()
            with self._lock:
                self._sweep_at = max(1024, 2 * len(self._store))

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

    def delete_for_job(self, job_key: str) -> int:
        """Forget every record of a job that no longer exists (`<job key>/<replica type>/pods|services`).  The
        reference keeps its records in a TTL store that drops them after five minutes; a plain dict never does, and a
        long-running operator kept three to four records per job it had ever seen."""
        prefix = job_key + "/"
        with self._lock:
            dead = [k for k in self._store if k.startswith(prefix)]
            for k in dead:
                del self._store[k]
        return len(dead)

    def sweep(self) -> int:
        """Drop the records that are both settled and older than the TTL (what the TTL store does by itself)."""
        now = self.clock()
        with self._lock:
            dead = [k for k, r in self._store.items() if now - r.ts > self.ttl]
            for k in dead:
                del self._store[k]
        return len(dead)

    def __len__(self) -> int:
        with self._lock:
            return len(self._store)


def gen_expectation_pods_key(job_key: str, rtype: str) -> str:
    return f"{job_key}/{rtype.lower()}/pods"


def gen_expectation_services_key(job_key: str, rtype: str) -> str:
    return f"{job_key}/{rtype.lower()}/services"
