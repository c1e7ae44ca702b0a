"""PyTorchJob controller (reference controllers/pytorch/: pytorchjob_controller.go:71-262,
status.go:34-146, job.go, pod.go, service.go).

Rendezvous contract injected into EVERY container of every pod (pytorchjob_controller.go:180-233):
  MASTER_PORT = the Master's `pytorchjob-port`      MASTER_ADDR = localhost (master) | <job>-master-0
  WORLD_SIZE  = Σ replicas                           RANK = 0 (master) | index+1 (worker)
  PYTHONUNBUFFERED = 0
This is exactly what kubedl_b200.worker.ddp_trainer.init_distributed consumes (env://).
"""
from __future__ import annotations

from typing import Dict, List

from ..api import common as c
from ..api import kinds as k
from ..engine.job_controller import gen_general_name, get_port_from_job
from ..utils import status as st
from .base import BaseJobReconciler


def append_env(container: dict, name: str, value: str) -> None:
    env = container.get("env")
    if env is None:
        env = []
        container["env"] = env
    env.append({"name": name, "value": str(value)})


def contain_master_spec(job: dict) -> bool:
    return k.PYTORCH_MASTER in (job.get("spec", {}).get("pytorchReplicaSpecs") or {})


class PytorchJobReconciler(BaseJobReconciler):
    KIND = k.PYTORCH_KIND
    CONTROLLER_NAME = "PytorchJobController"

    def set_cluster_spec(self, job: dict, pod_template: dict, rtype: str, index: str) -> None:
        specs = job["spec"]["pytorchReplicaSpecs"]
        rank = int(index)
        master_port = get_port_from_job(specs, k.PYTORCH_MASTER, k.PYTORCH_DEFAULT_CONTAINER,
                                        k.PYTORCH_DEFAULT_PORT_NAME)
        master_addr = gen_general_name(c.name_of(job), k.PYTORCH_MASTER.lower(), "0")
        if rtype == k.PYTORCH_MASTER.lower():
            if rank != 0:
                raise ValueError("invalid config: There should be only a single master with index=0")
            master_addr = "localhost"
        else:
            rank += 1
        total = c.total_replicas(specs)
        for ctr in c.containers_of(pod_template):
            append_env(ctr, "MASTER_PORT", master_port)
            append_env(ctr, "MASTER_ADDR", master_addr)
            append_env(ctr, "WORLD_SIZE", total)
            append_env(ctr, "RANK", rank)
            append_env(ctr, "PYTHONUNBUFFERED", "0")

    def get_reconcile_orders(self) -> List[str]:
        return [k.PYTORCH_MASTER, k.PYTORCH_WORKER]

    def is_master_role(self, replicas: Dict[str, dict], rtype: str, index: int) -> bool:
        return k.PYTORCH_MASTER in replicas and rtype == k.PYTORCH_MASTER

    def needs_service_for(self, rtype: str) -> bool:
        """Only the Master gets a headless Service (reference pkg/job_controller/job.go:223-227)."""
        return rtype == k.PYTORCH_MASTER

    def update_job_status(self, job: dict, replicas: Dict[str, dict], job_status: dict, restart: bool) -> None:
        """reference controllers/pytorch/status.go:34-125"""
        name = c.name_of(job)
        if not job_status.get("startTime"):
            job_status["startTime"] = c.now_rfc3339()
        prev_restarting, prev_failed = st.is_restarting(job_status), st.is_failed(job_status)
        for rtype, spec in replicas.items():
            status = (job_status.get("replicaStatuses") or {}).get(rtype)
            if status is None:
                continue
            expected = c.replicas_of(spec) - int(status.get("succeeded", 0))
            running, failed = int(status.get("active", 0)), int(status.get("failed", 0))
            if not contain_master_spec(job):
                raise ValueError("invalid config: Job must contain master replica spec")
            if rtype == k.PYTORCH_MASTER:
                if running > 0:
                    st.update_job_conditions(job_status, c.JOB_RUNNING, st.JOB_RUNNING_REASON,
                                             f"PyTorchJob {name} is running.")
                if expected == 0:
                    self._mark_succeeded(job, job_status, f"PyTorchJob {name} is successfully completed.")
            if failed > 0:
                self._mark_failed_or_restarting(job, job_status, rtype, failed, restart, prev_restarting, prev_failed)
