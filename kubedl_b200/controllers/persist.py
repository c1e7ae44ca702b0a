"""Persist controllers: mirror jobs, pods and events into the storage backends
(reference controllers/persist/: persist_controller.go:30-74, object/job/*_persist_controller.go,
object/pod/pod_persist_controller.go:81-140, event/event_persist_controller.go:74-103, custom
enqueue handlers object/job/job_event_handler.go:38-68, util/request.go:26-38,
filters util/filter.go:30-41, event/events_event_handler.go:86-108).

Requests carry "<uid>/<name>" in the name field so a delete can still address the stored row
after the object is gone.  Enabled by --object-storage / --event-storage; --region (or $REGION)
tags the rows.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Tuple

from ..api import common as c
from ..api.kinds import SCHEME
from ..apiserver.store import NotFound
from ..plugins.storage import backends as B
from ..plugins.storage import dmo
from ..runtime.manager import Manager, Predicates, Request, Result
from ..utils.logger import logger
from ..utils.workloadgate import AUTO, is_workload_enable

log = logger("kubedl.persist")


def gen_id_name(uid: str, name: str) -> str:
    return f"{uid}/{name}"


def parse_id_name(s: str) -> Tuple[str, str]:
    uid, _, name = s.partition("/")
    if not name:
        raise ValueError(f"bad persist request name {s!r}")
    return uid, name


def _enqueue_with_uid(ev, obj, old) -> List[Request]:
    return [Request(c.namespace_of(obj), gen_id_name(c.uid_of(obj), c.name_of(obj)))]


def is_kubedl_managed(obj: dict) -> bool:
    """Only children of one of the four job kinds are persisted (reference util/filter.go:30-41)."""
    ref = c.controller_ref(obj)
    return ref is not None and SCHEME.is_workload(ref.get("kind", ""))


class JobPersistController:
    def __init__(self, mgr: Manager, kind: str, backend: B.ObjectStorageBackend, region: str):
        self.mgr, self.kind, self.backend, self.region = mgr, kind, backend, region
        self.ctl = mgr.new_controller(f"{kind}PersistController", self.reconcile)
        self.ctl.watch(kind, _enqueue_with_uid)

    def reconcile(self, req: Request) -> Result:
        uid, name = parse_id_name(req.name)
        try:
            job = self.mgr.client.get(self.kind, req.namespace, name)
        except NotFound:
            # gone from the API server: mark stopped + deleted using the id carried by the request
            self.backend.stop_job(req.namespace, name, uid, self.region)
            self.backend.delete_job(req.namespace, name, uid, self.region)
            return Result()
        if c.uid_of(job) != uid:
            return Result()
        kind, specs, status = dmo.extract_typed_job_infos(job)
        self.backend.save_job(job, kind, specs, status, self.region)
        return Result()


class PodPersistController:
    def __init__(self, mgr: Manager, backend: B.ObjectStorageBackend, region: str):
        self.mgr, self.backend, self.region = mgr, backend, region
        self.ctl = mgr.new_controller("PodPersistController", self.reconcile)
        pred = Predicates(create=is_kubedl_managed, update=lambda old, new: is_kubedl_managed(new),
                          delete=is_kubedl_managed)
        self.ctl.watch("Pod", _enqueue_with_uid, pred)

    def reconcile(self, req: Request) -> Result:
        uid, name = parse_id_name(req.name)
        try:
            pod = self.mgr.client.get("Pod", req.namespace, name)
        except NotFound:
            self.backend.stop_pod(req.namespace, name, uid)
            return Result()
        ref = c.controller_ref(pod)
        default_container = SCHEME.info(ref["kind"]).default_container_name if ref and SCHEME.is_workload(ref["kind"]) else ""
        self.backend.save_pod(pod, default_container, self.region)
        return Result()


class EventPersistController:
    def __init__(self, mgr: Manager, backend: B.EventStorageBackend, region: str):
        self.mgr, self.backend, self.region = mgr, backend, region
        self.ctl = mgr.new_controller("EventPersistController", self.reconcile)
        self.ctl.watch("Event", lambda ev, obj, old: [Request(c.namespace_of(obj), c.name_of(obj))],
                       Predicates(create=self._wanted, update=lambda o, n: self._wanted(n), delete=lambda o: False))

    @staticmethod
    def _wanted(event: dict) -> bool:
        """Events about the four job kinds, Pods and Services (reference events_event_handler.go:86-108)."""
        kind = (event.get("involvedObject") or {}).get("kind", "")
        return SCHEME.is_workload(kind) or kind in ("Pod", "Service")

    def reconcile(self, req: Request) -> Result:
        try:
            ev = self.mgr.client.get("Event", req.namespace, req.name)
        except NotFound:
            return Result()
        self.backend.save_event(ev, self.region)
        return Result()


def setup_with_manager(mgr: Manager, object_storage: str = "", event_storage: str = "", region: str = "",
                       workloads: str = AUTO) -> Dict[str, object]:
    """reference persist_controller.go:42-74"""
    region = region or os.environ.get("REGION", "")
    out: Dict[str, object] = {}
    if object_storage:
        backend = B.REGISTRY.get_object_backend(object_storage)
        if backend is None:
            raise ValueError(f"object storage backend [{object_storage}] has not registered")
        backend.initialize()
        for info in SCHEME.workload_kinds():
            if is_workload_enable(info.kind, workloads, crd_installed=mgr.api.crd_installed):
                out[f"job/{info.kind}"] = JobPersistController(mgr, info.kind, backend, region)
        out["pod"] = PodPersistController(mgr, backend, region)
    if event_storage:
        backend = B.REGISTRY.get_event_backend(event_storage)
        if backend is None:
            raise ValueError(f"event storage backend [{event_storage}] has not registered")
        backend.initialize()
        out["event"] = EventPersistController(mgr, backend, region)
    return out
