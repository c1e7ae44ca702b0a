"""XGBoostJob controller (reference controllers/xgboost/: xgboostjob_controller.go:65-200,
pod.go:106-152, job.go:89-189, service.go:42-67).

Env on every container: MASTER_PORT, MASTER_ADDR=<job>-master-0 (for the master too),
WORLD_SIZE, RANK=<replica index>, PYTHONUNBUFFERED=0.  RANK is the bare index for BOTH replica
types, so master-0 and worker-0 share RANK=0 (encoded in the reference test pod_test.go:84-137).
Parity is kept here; the worker runtime derives a unique global rank itself
(kubedl_b200.worker.rendezvous.global_rank_from_env).  Other quirks kept: CreateService
swallows errors (service.go:42-45), DeleteService ignores errors once the job succeeded
(:59-67), a Running condition is re-asserted at the end of every non-terminal pass
(job.go:180-188).
"""
from __future__ import annotations

from typing import Dict, List

from ..api import common as c
from ..api import kinds as k
from ..engine.job_controller import gen_general_name, get_port_from_job
from ..utils import status as st
from .base import BaseJobReconciler
from .pytorch import append_env


def set_pod_env(job: dict, pod_template: dict, index: str) -> None:
    specs = job["spec"]["xgbReplicaSpecs"]
    rank = int(index)
    master_addr = gen_general_name(c.name_of(job), k.XGB_MASTER.lower(), "0")
    master_port = get_port_from_job(specs, k.XGB_MASTER, k.XGB_DEFAULT_CONTAINER, k.XGB_DEFAULT_PORT_NAME)
    total = c.total_replicas(specs)
    for ctr in c.containers_of(pod_template):
        append_env(ctr, "MASTER_PORT", master_port)
        append_env(ctr, "MASTER_ADDR", master_addr)
        append_env(ctr, "WORLD_SIZE", total)
        append_env(ctr, "RANK", rank)
        append_env(ctr, "PYTHONUNBUFFERED", "0")


class XgboostJobReconciler(BaseJobReconciler):
    KIND = k.XGB_KIND
    CONTROLLER_NAME = "XGBoostJobController"

    def set_cluster_spec(self, job: dict, pod_template: dict, rtype: str, index: str) -> None:
        set_pod_env(job, pod_template, index)

    def get_reconcile_orders(self) -> List[str]:
        return [k.XGB_MASTER, k.XGB_WORKER]

    def is_master_role(self, replicas, rtype: str, index: int) -> bool:
        return k.XGB_MASTER in replicas and rtype == k.XGB_MASTER

    def create_service(self, job: dict, service: dict) -> None:
        try:
            super().create_service(job, service)
        except Exception as e:  # logged and swallowed upstream
            self.log.warning("create service %s failed: %s", c.name_of(service), e)

    def delete_service(self, job: dict, name: str, namespace: str) -> None:
        try:
            super().delete_service(job, name, namespace)
        except Exception:
            if st.is_succeeded(c.status_of(job)):
                return
            raise

    def update_job_status(self, job: dict, replicas: Dict[str, dict], job_status: dict, restart: bool) -> None:
        name = c.name_of(job)
        prev_restarting, prev_failed = st.is_restarting(job_status), st.is_failed(job_status)
        for rtype, spec in replicas.items():
            status = (job_status.get("replicaStatuses") or {}).get(rtype)
            if status is None:
                continue
            succeeded = int(status.get("succeeded", 0))
            expected = c.replicas_of(spec) - succeeded
            running, failed = int(status.get("active", 0)), int(status.get("failed", 0))
            if running == c.replicas_of(spec) and not job_status.get("startTime"):
                job_status["startTime"] = c.now_rfc3339()
            if rtype == k.XGB_MASTER:
                if running > 0:
                    st.update_job_conditions(job_status, c.JOB_RUNNING, st.JOB_RUNNING_REASON,
                                             f"XGBoostJob {name} is running.")
                if expected == 0:
                    self._mark_succeeded(job, job_status, f"XGBoostJob {name} is successfully completed.")
                    return
            if failed > 0:
                self._mark_failed_or_restarting(job, job_status, rtype, failed, restart, prev_restarting, prev_failed)
        st.update_job_conditions(job_status, c.JOB_RUNNING, st.JOB_RUNNING_REASON, f"XGBoostJob {name} is running.")
