"""TFJob controller (reference controllers/tensorflow/: tfjob_controller.go:75-276,
tensorflow.go:40-142, status.go:56-212).

`TF_CONFIG` is injected only into the container named `tensorflow` and only for distributed
jobs (Σ replicas != 1).  Cluster hosts are `<job>-<rt>-<i>.<ns>.svc[.$CUSTOM_CLUSTER_DOMAIN]:<port>`;
Evaluator replicas are left out of the cluster map.  JSON layout matches Go's encoder:
{"cluster":{<sorted types>},"task":{"type":..,"index":..},"environment":"cloud"}.
Quirk kept: `Evaluator` is accepted and counted but never reconciled because it is missing from
the reconcile order (tfjob_controller.go:263-270).
"""
from __future__ import annotations

import json
import os
from typing import Dict, List

from ..api import common as c
from ..api import kinds as k
from ..engine.job_controller import gen_general_name, get_port_from_job
from ..utils import k8sutil
from ..utils import status as st
from .base import BaseJobReconciler
from .pytorch import append_env

ENV_CUSTOM_CLUSTER_DOMAIN = "CUSTOM_CLUSTER_DOMAIN"
TF_CONFIG = "TF_CONFIG"


def is_distributed(job: dict) -> bool:
    specs = job.get("spec", {}).get("tfReplicaSpecs") or {}
    count = 0
    for t in (k.TF_CHIEF, k.TF_EVAL, k.TF_MASTER, k.TF_PS, k.TF_WORKER):
        if specs.get(t) is not None:
            count += c.replicas_of(specs[t])
    return count != 1


def gen_cluster_spec(job: dict) -> Dict[str, List[str]]:
    specs = job["spec"]["tfReplicaSpecs"]
    cluster: Dict[str, List[str]] = {}
    domain = os.environ.get(ENV_CUSTOM_CLUSTER_DOMAIN, "")
    for rtype, spec in specs.items():
        if rtype == k.TF_EVAL:
            continue  # evaluators are not part of the training cluster
        rt = rtype.lower()
        port = get_port_from_job(specs, rtype, k.TF_DEFAULT_CONTAINER, k.TF_DEFAULT_PORT_NAME)
        hosts = []
        for i in range(c.replicas_of(spec)):
            svc = f"{gen_general_name(c.name_of(job), rt, str(i))}.{c.namespace_of(job)}.svc"
            if domain:
                svc += "." + domain
            hosts.append(f"{svc}:{port}")
        cluster[rt] = hosts
    return cluster


def gen_tf_config_json(job: dict, rtype: str, index: str) -> str:
    cfg = {"cluster": dict(sorted(gen_cluster_spec(job).items())),
           "task": {"type": rtype, "index": int(index)}, "environment": "cloud"}
    return json.dumps(cfg, separators=(",", ":"))


def contain_chief_or_master_spec(job: dict) -> bool:
    specs = job.get("spec", {}).get("tfReplicaSpecs") or {}
    return k.TF_CHIEF in specs or k.TF_MASTER in specs


class TFJobReconciler(BaseJobReconciler):
    KIND = k.TF_KIND
    CONTROLLER_NAME = "TFJobController"

    def set_cluster_spec(self, job: dict, pod_template: dict, rtype: str, index: str) -> None:
        if not is_distributed(job):
            return
        cfg = gen_tf_config_json(job, rtype, index)
        for ctr in c.containers_of(pod_template):
            if ctr.get("name") == k.TF_DEFAULT_CONTAINER:
                append_env(ctr, TF_CONFIG, cfg)
                break

    def get_reconcile_orders(self) -> List[str]:
        return [k.TF_PS, k.TF_MASTER, k.TF_CHIEF, k.TF_WORKER]

    def is_master_role(self, replicas, rtype: str, index: int) -> bool:
        return k.tf_is_chief_or_master(rtype)

    def update_job_status(self, job: dict, replicas: Dict[str, dict], job_status: dict, restart: bool) -> None:
        """reference controllers/tensorflow/status.go:56-212"""
        name = c.name_of(job)
        prev_restarting, prev_failed = st.is_restarting(job_status), st.is_failed(job_status)
        worker0_completed = False
        pods = self.ctrl.filter_pods_for_replica_type(self.get_pods_for_job(job), k.TF_WORKER.lower())
        for pod in pods:
            try:
                idx = int(c.labels_of(pod).get(c.REPLICA_INDEX_LABEL, ""))
            except ValueError:
                continue
            if idx == 0:
                exit_code = 0xBEEF
                for cs in (pod.get("status") or {}).get("containerStatuses") or []:
                    term = (cs.get("state") or {}).get("terminated")
                    if cs.get("name") == k.TF_DEFAULT_CONTAINER and term is not None:
                        exit_code = int(term.get("exitCode", 0))
                        break
                if exit_code == 0 and k8sutil.phase_of(pod) == c.POD_SUCCEEDED:
                    worker0_completed = True
                break
        if not job_status.get("startTime"):
            job_status["startTime"] = c.now_rfc3339()
        for rtype, spec in replicas.items():
            status = (job_status.get("replicaStatuses") or {}).get(rtype)
            if status is None:
                continue
            expected = c.replicas_of(spec) - int(status.get("succeeded", 0))
            running, failed = int(status.get("active", 0)), int(status.get("failed", 0))
            if contain_chief_or_master_spec(job):
                if k.tf_is_chief_or_master(rtype):
                    if running > 0:
                        st.update_job_conditions(job_status, c.JOB_RUNNING, st.JOB_RUNNING_REASON,
                                                 f"TFJob {name} is running.")
                    if expected == 0:
                        self._mark_succeeded(job, job_status, f"TFJob {name} successfully completed.")
            elif rtype == k.TF_WORKER:
                if expected == 0 or worker0_completed:
                    self._mark_succeeded(job, job_status, f"TFJob {name} successfully completed.")
                elif running > 0:
                    st.update_job_conditions(job_status, c.JOB_RUNNING, st.JOB_RUNNING_REASON,
                                             f"TFJob {name} is running.")
            if failed > 0:
                self._mark_failed_or_restarting(job, job_status, rtype, failed, restart, prev_restarting, prev_failed)
