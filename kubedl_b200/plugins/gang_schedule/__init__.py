"""Gang scheduling plugin interface + registry (reference pkg/gang_schedule/interface.go:30-52,
registry/registry.go:27-73, registry/add_batch_scheduler.go:23-25).

Two implementations are registered:
  kube-batch   the reference's only one — a PodGroup named after the job with
               MinMember = Σ replicas, bind = set schedulerName (batch_scheduler/scheduler.go:44-119)
  gpu-gang     single-node all-or-nothing GPU placement for the 8×B200 box: the PodGroup also
               records the GPU demand, and the node agent admits the job's pods only when
               MinMember GPUs are free at once (SURVEY.md §2.3 last row).
"""
from __future__ import annotations

import abc
import threading
from typing import Callable, Dict, List, Optional

from ...api import common as c
from ...apiserver.store import NotFound, is_not_found

PODGROUP_KIND = "PodGroup"
PODGROUP_API_VERSION = "scheduling.incubator.k8s.io/v1alpha1"


class GangScheduler(abc.ABC):
    @abc.abstractmethod
    def create_gang(self, job: dict, replicas: Dict[str, dict]) -> dict: ...
    @abc.abstractmethod
    def bind_pod_to_gang(self, pod_template: dict, gang: dict) -> None: ...
    @abc.abstractmethod
    def get_gang(self, namespace: str, name: str) -> dict: ...
    @abc.abstractmethod
    def delete_gang(self, namespace: str, name: str) -> None: ...
    @abc.abstractmethod
    def name(self) -> str: ...


def append_owner_reference(obj: dict, ref: dict) -> None:
    """reference pkg/gang_schedule/utils.go:23-41: no duplicates by UID."""
    refs = obj.setdefault("metadata", {}).setdefault("ownerReferences", [])
    if not any(r.get("uid") == ref.get("uid") for r in refs):
        refs.append(ref)


class KubeBatchScheduler(GangScheduler):
    SCHEDULER_NAME = "kube-batch"

    def __init__(self, client):
        self.client = client

    def name(self) -> str:
        return self.SCHEDULER_NAME

    def _pod_group(self, job: dict, replicas: Dict[str, dict]) -> dict:
        pg = {"apiVersion": PODGROUP_API_VERSION, "kind": PODGROUP_KIND,
              "metadata": {"name": c.name_of(job), "namespace": c.namespace_of(job)},
              "spec": {"minMember": c.total_replicas(replicas)}}
        append_owner_reference(pg, {"apiVersion": job.get("apiVersion", ""), "kind": job.get("kind", ""),
                                    "name": c.name_of(job), "uid": c.uid_of(job),
                                    "blockOwnerDeletion": True, "controller": True})
        return pg

    def create_gang(self, job: dict, replicas: Dict[str, dict]) -> dict:
        return self.client.create(self._pod_group(job, replicas))

    def bind_pod_to_gang(self, pod_template: dict, gang: dict) -> None:
        pod_template.setdefault("spec", {})["schedulerName"] = self.name()

    def get_gang(self, namespace: str, name: str) -> dict:
        return self.client.reader().get(PODGROUP_KIND, namespace, name)

    def delete_gang(self, namespace: str, name: str) -> None:
        try:
            self.client.delete(PODGROUP_KIND, namespace, name)
        except NotFound:
            return


class GpuGangScheduler(KubeBatchScheduler):
    """All-or-nothing GPU admission on one node.  Pods are annotated with their gang so the node
    agent can hold them Pending until the whole gang fits."""
    SCHEDULER_NAME = "gpu-gang"
    GANG_ANNOTATION = "kubedl.io/gang-name"

    def _pod_group(self, job, replicas):
        pg = super()._pod_group(job, replicas)
        gpus = 0
        for spec in replicas.values():
            per_pod = 0
            for ctr in spec.get("template", {}).get("spec", {}).get("containers") or []:
                lim = (ctr.get("resources") or {}).get("limits") or {}
                per_pod += int(lim.get("nvidia.com/gpu", 0) or 0)
            gpus += per_pod * c.replicas_of(spec)
        pg["spec"]["minResources"] = {"nvidia.com/gpu": gpus}
        # RunPolicy.schedulingPolicy.minAvailable is declared but unused by the reference
        # (SURVEY.md §2.6 item 18); here it lowers the gang size when present
        min_avail = ((c.run_policy_of(job).get("schedulingPolicy") or {}).get("minAvailable"))
        if min_avail is not None and 0 < int(min_avail) <= pg["spec"]["minMember"]:
            pg["spec"]["minMember"] = int(min_avail)
        return pg

    def bind_pod_to_gang(self, pod_template, gang):
        super().bind_pod_to_gang(pod_template, gang)
        ann = pod_template.setdefault("metadata", {}).setdefault("annotations", {})
        ann[self.GANG_ANNOTATION] = c.name_of(gang)


# ---------------------------------------------------------------------- registry
NewGangScheduler = Callable[[object], GangScheduler]   # factory(manager) -> scheduler

NEW_GANG_SCHEDULERS: List[NewGangScheduler] = [
    lambda mgr: KubeBatchScheduler(mgr.client),
    lambda mgr: GpuGangScheduler(mgr.client),
]


class Registry:
    def __init__(self):
        self._lock = threading.Lock()
        self._by_name: Dict[str, GangScheduler] = {}

    def add(self, gs: GangScheduler) -> None:
        with self._lock:
            self._by_name[gs.name()] = gs

    def get(self, name: str) -> Optional[GangScheduler]:
        with self._lock:
            return self._by_name.get(name)

    def remove(self, name: str) -> None:
        with self._lock:
            self._by_name.pop(name, None)

    def names(self) -> List[str]:
        with self._lock:
            return sorted(self._by_name)


REGISTRY = Registry()


def register_gang_schedulers(mgr) -> None:
    """reference registry.go:32-38 (called from main.go:88)."""
    for factory in NEW_GANG_SCHEDULERS:
        REGISTRY.add(factory(mgr))


def get(name: str) -> Optional[GangScheduler]:
    return REGISTRY.get(name)
