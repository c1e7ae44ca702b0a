"""Database model objects (reference pkg/storage/dmo/types.go:30-168): rows of `replica_info`,
`job_info` and `event_info`, and the converters from Kubernetes objects
(reference pkg/storage/dmo/converters/{job,pod,event}.go).

Reference bugs deliberately NOT reproduced: `Event.Count` no longer
shares the `reason` column, and the involved-object fields of an Event are carried over.
"""
from __future__ import annotations

import json
from dataclasses import asdict, dataclass, field
from typing import Dict, Optional

from ...api import common as c
from ...utils import k8sutil, quota, tenancy


class NoDependentOwner(ValueError):
    def __init__(self):
        super().__init__("object has no dependent owner")


class NoReplicaTypeLabel(ValueError):
    def __init__(self):
        super().__init__(f"object has no replica type label [{c.REPLICA_TYPE_LABEL}]")


@dataclass
class Pod:
    name: str = ""
    namespace: str = ""
    pod_id: str = ""
    version: str = ""
    status: str = ""
    image: str = ""
    job_id: str = ""
    replica_type: str = ""
    resources: str = ""
    host_ip: Optional[str] = None
    pod_ip: Optional[str] = None
    deploy_region: Optional[str] = None
    deleted: Optional[int] = None
    is_in_etcd: Optional[int] = None
    remark: Optional[str] = None
    gmt_created: Optional[str] = None
    gmt_modified: Optional[str] = None
    gmt_started: Optional[str] = None
    gmt_finished: Optional[str] = None
    TABLE = "replica_info"


@dataclass
class Job:
    name: str = ""
    namespace: str = ""
    job_id: str = ""
    version: str = ""
    status: str = ""
    kind: str = ""
    resources: str = ""
    deploy_region: Optional[str] = None
    tenant: Optional[str] = None
    owner: Optional[str] = None
    deleted: Optional[int] = None
    is_in_etcd: Optional[int] = None
    gmt_created: Optional[str] = None
    gmt_modified: Optional[str] = None
    gmt_finished: Optional[str] = None
    TABLE = "job_info"


@dataclass
class Event:
    name: str = ""
    kind: str = ""
    type: str = ""
    obj_namespace: str = ""
    obj_name: str = ""
    obj_uid: str = ""
    reason: str = ""
    message: str = ""
    count: int = 0
    region: Optional[str] = None
    first_timestamp: Optional[str] = None
    last_timestamp: Optional[str] = None
    TABLE = "event_info"


def _wire(res) -> dict:
    w = quota.resources_to_wire(res)
    return {k: v for k, v in w.items() if v}


def compute_pod_resources(pod_spec: dict) -> dict:
    """max(max over init containers, sum over containers) (reference converters/pod.go:147-154)."""
    return _wire(quota.pod_resources(pod_spec or {}))


def compute_job_resources(specs: Dict[str, dict]) -> dict:
    out = {}
    for rtype, spec in (specs or {}).items():
        rr = {"resources": compute_pod_resources(spec.get("template", {}).get("spec", {})), "replicas": 0}
        if spec.get("replicas") is not None:
            rr["replicas"] = int(spec["replicas"])
        out[rtype] = rr
    return out


def convert_job_to_dmo_job(job: dict, kind: str, specs: Dict[str, dict], job_status: dict, region: str = "") -> Job:
    m = job.get("metadata", {})
    d = Job(name=c.name_of(job), namespace=c.namespace_of(job), job_id=c.uid_of(job),
            version=m.get("resourceVersion", ""), kind=kind, gmt_created=m.get("creationTimestamp"))
    if region:
        d.deploy_region = region
    tn = None
    try:
        tn = tenancy.get_tenancy(job)
    except ValueError:
        tn = None
    if tn is not None:
        d.tenant, d.owner = tn.tenant, tn.user
        if d.deploy_region is None and tn.region:
            d.deploy_region = tn.region
    else:
        d.tenant, d.owner = "", ""
    conds = (job_status or {}).get("conditions") or []
    d.status = conds[-1].get("type", c.JOB_CREATED) if conds else c.JOB_CREATED
    if (job_status or {}).get("completionTime"):
        d.gmt_finished = job_status["completionTime"]
    d.deleted, d.is_in_etcd = 0, 1
    d.resources = json.dumps(compute_job_resources(specs), sort_keys=True)
    return d


def _pick(items, default_name: str, key="name"):
    chosen = items[0]
    for it in items[1:]:
        if it.get(key) == default_name:
            chosen = it
            break
    return chosen


def convert_pod_to_dmo_pod(pod: dict, default_container_name: str, region: str = "") -> Pod:
    m = pod.get("metadata", {})
    d = Pod(name=c.name_of(pod), namespace=c.namespace_of(pod), pod_id=c.uid_of(pod),
            version=m.get("resourceVersion", ""), gmt_created=m.get("creationTimestamp"))
    if region:
        d.deploy_region = region
    owner_id, _ = k8sutil.resolve_dependent_owner(pod)
    if not owner_id:
        raise NoDependentOwner()
    d.job_id = owner_id
    rtype, ok = k8sutil.get_replica_type(pod)
    if not ok:
        raise NoReplicaTypeLabel()
    d.replica_type = rtype
    d.resources = json.dumps(compute_pod_resources(pod.get("spec", {})), sort_keys=True)
    d.deleted, d.is_in_etcd = 0, 1
    status = pod.get("status") or {}
    if status.get("podIP"):
        d.pod_ip = status["podIP"]
    if status.get("hostIP"):
        d.host_ip = status["hostIP"]
    containers = pod.get("spec", {}).get("containers") or []
    if not containers:
        return d
    d.image = _pick(containers, default_container_name).get("image", "")
    d.status = c.POD_UNKNOWN
    cstats = status.get("containerStatuses") or []
    if not cstats:
        return d
    cs = _pick(cstats, default_container_name)
    phase = status.get("phase", "")
    d.status = phase
    state = cs.get("state") or {}
    created = m.get("creationTimestamp")
    if phase == c.POD_RUNNING:
        if state.get("running"):
            d.gmt_started = state["running"].get("startedAt")
        if not d.gmt_started:
            d.gmt_started = created
    elif phase in (c.POD_SUCCEEDED, c.POD_FAILED):
        term = state.get("terminated")
        if term:
            d.gmt_started = term.get("startedAt")
            d.gmt_finished = term.get("finishedAt")
            if phase == c.POD_FAILED:
                d.remark = (f"Reason: {term.get('reason', '')}\nExitCode: {term.get('exitCode', 0)}\n"
                            f"Message: {term.get('message', '')}")
        if not d.gmt_started:
            d.gmt_started = created
        if not d.gmt_finished:
            d.gmt_finished = c.now_rfc3339()
    return d


def convert_event_to_dmo_event(event: dict, region: str = "") -> Event:
    inv = event.get("involvedObject") or {}
    d = Event(name=c.name_of(event), kind=inv.get("kind", ""), type=event.get("type", ""),
              obj_namespace=inv.get("namespace", ""), obj_name=inv.get("name", ""), obj_uid=inv.get("uid", ""),
              reason=event.get("reason", ""), message=event.get("message", ""), count=int(event.get("count", 0) or 0),
              first_timestamp=event.get("firstTimestamp"), last_timestamp=event.get("lastTimestamp"))
    if region:
        d.region = region
    return d


def extract_typed_job_infos(job: dict):
    """(kind, replica specs, status) of any of the four kinds (reference converters/job.go:87-100)."""
    from ...api.kinds import SCHEME
    kind = job.get("kind", "")
    if not SCHEME.is_workload(kind):
        raise ValueError(f"unknown job kind, {c.namespace_of(job)}/{c.name_of(job)}")
    return kind, SCHEME.info(kind).replica_specs(job), c.status_of(job)
