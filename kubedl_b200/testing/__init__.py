"""Test fixtures for people extending the framework (reference pkg/test_job/v1/,
pkg/test_util/v1/, pkg/job_controller/test_job_controller.go): a `TestJob` kind, a
`TestJobController` stub that implements ControllerInterface over in-memory lists, and the
builders NewTestJob / NewPod / NewPodList / NewBaseService / SetPodsStatuses / CheckCondition.
"""
from __future__ import annotations

from typing import Dict, List, Optional

from ..api import common as c
from ..api.kinds import SCHEME, KindInfo
from ..engine.interface import ControllerInterface
from ..utils import status as st

TEST_JOB_KIND = "TestJob"
TEST_GROUP = "kubedl.io"
TEST_VERSION = "v1"
TEST_REPLICA_WORKER = "Worker"
TEST_REPLICA_MASTER = "Master"
TEST_IMAGE = "test-image-for-kubedl:latest"
DEFAULT_CONTAINER = "test-container"
DEFAULT_PORT_NAME = "test-port"
DEFAULT_PORT = 2222
LABEL_GROUP_NAME = "test.kubedl.io"


def register_test_kind() -> KindInfo:
    if not SCHEME.is_workload(TEST_JOB_KIND):
        SCHEME.register(KindInfo(TEST_JOB_KIND, TEST_GROUP, TEST_VERSION, "testjobs", "testjob", "testReplicaSpecs",
                                 (TEST_REPLICA_MASTER, TEST_REPLICA_WORKER), LABEL_GROUP_NAME, DEFAULT_CONTAINER,
                                 DEFAULT_PORT_NAME, DEFAULT_PORT, lambda job: None))
    return SCHEME.info(TEST_JOB_KIND)


def new_test_replica_spec_template() -> dict:
    return {"spec": {"containers": [{"name": DEFAULT_CONTAINER, "image": TEST_IMAGE, "args": ["Fake", "Fake"],
                                     "ports": [{"name": DEFAULT_PORT_NAME, "containerPort": DEFAULT_PORT}]}]}}


def new_test_job(workers: int, name: str = "test-job", namespace: str = "default") -> dict:
    job = {"apiVersion": f"{TEST_GROUP}/{TEST_VERSION}", "kind": TEST_JOB_KIND,
           "metadata": {"name": name, "namespace": namespace, "uid": f"uid-{name}"},
           "spec": {"testReplicaSpecs": {}}, "status": c.new_job_status()}
    if workers > 0:
        job["spec"]["testReplicaSpecs"][TEST_REPLICA_WORKER] = {"replicas": workers, "template": new_test_replica_spec_template()}
    return job


def gen_labels(job_name: str) -> Dict[str, str]:
    return {c.GROUP_NAME_LABEL: LABEL_GROUP_NAME, c.JOB_NAME_LABEL: job_name.replace("/", "-")}


def gen_owner_reference(job: dict) -> dict:
    return {"apiVersion": f"{TEST_GROUP}/{TEST_VERSION}", "kind": TEST_JOB_KIND, "name": c.name_of(job),
            "uid": c.uid_of(job), "blockOwnerDeletion": True, "controller": True}


def new_base_pod(name: str, job: dict) -> dict:
    return {"apiVersion": "v1", "kind": "Pod",
            "metadata": {"name": name, "namespace": c.namespace_of(job), "labels": gen_labels(c.name_of(job)),
                         "ownerReferences": [gen_owner_reference(job)]},
            "spec": {"containers": []}, "status": {}}


def new_pod(job: dict, typ: str, index: int) -> dict:
    pod = new_base_pod(f"{typ}-{index}", job)
    pod["metadata"]["labels"][c.REPLICA_TYPE_LABEL] = typ
    pod["metadata"]["labels"][c.REPLICA_INDEX_LABEL] = str(index)
    return pod


def new_pod_list(count: int, phase: str, job: dict, typ: str, start: int = 0) -> List[dict]:
    pods = []
    for i in range(count):
        p = new_pod(job, typ, start + i)
        p["status"] = {"phase": phase}
        pods.append(p)
    return pods


def set_pods_statuses(store: List[dict], job: dict, typ: str, pending: int, active: int, succeeded: int,
                      failed: int, restart_counts: Optional[List[int]] = None) -> None:
    idx = 0
    for count, phase in ((pending, c.POD_PENDING), (active, c.POD_RUNNING), (succeeded, c.POD_SUCCEEDED), (failed, c.POD_FAILED)):
        pods = new_pod_list(count, phase, job, typ, idx)
        if phase == c.POD_RUNNING and restart_counts:
            for p, rc in zip(pods, restart_counts):
                p["status"]["containerStatuses"] = [{"restartCount": rc}]
        store.extend(pods)
        idx += count


def new_base_service(name: str, job: dict) -> dict:
    return {"apiVersion": "v1", "kind": "Service",
            "metadata": {"name": name, "namespace": c.namespace_of(job), "labels": gen_labels(c.name_of(job)),
                         "ownerReferences": [gen_owner_reference(job)]}, "spec": {}}


def new_service(job: dict, typ: str, index: int) -> dict:
    svc = new_base_service(f"{typ}-{index}", job)
    svc["metadata"]["labels"][c.REPLICA_TYPE_LABEL] = typ
    svc["metadata"]["labels"][c.REPLICA_INDEX_LABEL] = str(index)
    return svc


def check_condition(job: dict, cond_type: str, reason: str) -> bool:
    cond = st.get_condition(c.status_of(job), cond_type)
    return cond is not None and cond.get("status") == c.CONDITION_TRUE and cond.get("reason") == reason


class TestJobController(ControllerInterface):
    """ControllerInterface over in-memory slices (reference pkg/test_job/v1/test_job_controller.go:13-127)."""
    __test__ = False

    def __init__(self):
        self.job: Optional[dict] = None
        self.pods: List[dict] = []
        self.services: List[dict] = []

    def controller_name(self): return "test-operator"
    def get_api_group_version_kind(self): return (TEST_GROUP, TEST_VERSION, TEST_JOB_KIND)
    def get_api_group_version(self): return (TEST_GROUP, TEST_VERSION)
    def get_group_name_label_value(self): return LABEL_GROUP_NAME
    def get_job_from_informer_cache(self, namespace, name): return self.job
    def get_job_from_api_client(self, namespace, name): return self.job
    def delete_job(self, job): self.job = None
    def update_job_status(self, job, replicas, job_status, restart): return None
    def update_job_status_in_api_server(self, job, job_status): job["status"] = job_status
    def get_pods_for_job(self, job): return list(self.pods)
    def get_services_for_job(self, job): return list(self.services)
    def create_pod(self, job, pod): self.pods.append(pod)
    def create_service(self, job, service): self.services.append(service)

    def delete_pod(self, job, pod):
        self.pods = [p for p in self.pods if not (c.name_of(p) == c.name_of(pod) and c.namespace_of(p) == c.namespace_of(pod))]

    def delete_service(self, job, name, namespace):
        self.services = [s for s in self.services if not (c.name_of(s) == name and c.namespace_of(s) == namespace)]

    def set_cluster_spec(self, job, pod_template, rtype, index): return None
    def get_reconcile_orders(self): return [TEST_REPLICA_MASTER, TEST_REPLICA_WORKER]
    def is_master_role(self, replicas, rtype, index): return False
    def get_default_container_name(self): return DEFAULT_CONTAINER
    def get_default_container_port_name(self): return DEFAULT_PORT_NAME
    def get_default_container_port_number(self): return DEFAULT_PORT
