"""XDLJob controller (reference controllers/xdl/: xdljob_controller.go:79-248, status.go:61-160,
xdl.go:60-101).

Env on every container: TASK_NAME=<replica type, lower case>, TASK_INDEX=<index>; an existing
`ZK_ADDR` value gets `/<job UID>` appended so concurrent jobs do not collide in ZooKeeper.
Success rule: Σ succeeded Worker(+ExtendRole) >= minFinish, where minFinishWorkRate (percent,
ceil) wins over minFinishWorkNum and the default is all workers (status.go:151-160).
`gen_xdl_config` mirrors the reference's XDL_CONFIG generator, which exists but is never called.
"""
from __future__ import annotations

import json
import math
from typing import Dict, List

from ..api import common as c
from ..api import kinds as k
from ..engine.job_controller import gen_general_name, get_port_from_job
from ..utils import status as st
from .base import BaseJobReconciler
from .pytorch import append_env

ZK_ADDR, TASK_NAME, TASK_INDEX = "ZK_ADDR", "TASK_NAME", "TASK_INDEX"
XDL_JOB_FAILED_REASON = "XdlJobFailed"
XDL_JOB_RESTARTING_REASON = "XdlJobRestarting"


def calculate_min_finish(job: dict, workers: int) -> int:
    spec = job.get("spec", {})
    if spec.get("minFinishWorkRate") is not None:
        return int(math.ceil(workers * int(spec["minFinishWorkRate"]) / 100.0))
    if spec.get("minFinishWorkNum") is not None:
        return int(spec["minFinishWorkNum"])
    return workers


def gen_xdl_config(job: dict, rtype: str, index: str) -> str:
    """Cluster description {"cluster": {rt: ["host:port", ...]}, "task": {...}} (dead code
    upstream, reference controllers/xdl/xdl.go:60-101; offered to workers that want it)."""
    specs = job["spec"]["xdlReplicaSpecs"]
    cluster = {}
    for rt, spec in specs.items():
        port = get_port_from_job(specs, rt, k.XDL_DEFAULT_CONTAINER, k.XDL_DEFAULT_PORT_NAME)
        cluster[rt.lower()] = [f"{gen_general_name(c.name_of(job), rt.lower(), str(i))}:{port}"
                               for i in range(c.replicas_of(spec))]
    return json.dumps({"cluster": dict(sorted(cluster.items())), "task": {"type": rtype, "index": int(index)}},
                      separators=(",", ":"))


class XDLJobReconciler(BaseJobReconciler):
    KIND = k.XDL_KIND
    CONTROLLER_NAME = "XDLJobController"

    def set_cluster_spec(self, job: dict, pod_template: dict, rtype: str, index: str) -> None:
        uid = c.uid_of(job)
        for ctr in c.containers_of(pod_template):
            for env in ctr.get("env") or []:
                if env.get("name") == ZK_ADDR:
                    v = env.get("value", "")
                    env["value"] = v + uid if v.endswith("/") else v + "/" + uid
            append_env(ctr, TASK_NAME, rtype.lower())
            append_env(ctr, TASK_INDEX, index)

    def get_reconcile_orders(self) -> List[str]:
        return [k.XDL_PS, k.XDL_SCHEDULER, k.XDL_WORKER, k.XDL_EXTEND_ROLE]

    def is_master_role(self, replicas, rtype: str, index: int) -> bool:
        return False

    def update_job_status(self, job: dict, replicas: Dict[str, dict], job_status: dict, restart: bool) -> None:
        name = c.name_of(job)
        prev_restarting, prev_failed = st.is_restarting(job_status), st.is_failed(job_status)
        worker_num = worker_succeeded = 0
        for rtype, spec in replicas.items():
            status = (job_status.get("replicaStatuses") or {}).get(rtype)
            if status is None:
                continue
            n = c.replicas_of(spec)
            failed = int(status.get("failed", 0))
            if rtype in (k.XDL_WORKER, k.XDL_EXTEND_ROLE):
                worker_num += n
                worker_succeeded += int(status.get("succeeded", 0))
            if int(status.get("active", 0)) == n and not job_status.get("startTime"):
                job_status["startTime"] = c.now_rfc3339()
            if failed > 0:
                self._mark_failed_or_restarting(job, job_status, rtype, failed, restart, prev_restarting,
                                                prev_failed, restarting_reason=XDL_JOB_RESTARTING_REASON,
                                                failed_reason=XDL_JOB_FAILED_REASON)
                return
        if worker_succeeded >= calculate_min_finish(job, worker_num):
            self._mark_succeeded(job, job_status, f"XDLJob {name} is successfully completed.", emit_event=False)
            return
        st.update_job_conditions(job_status, c.JOB_RUNNING, st.JOB_RUNNING_REASON, f"XDLJob {name} is running.")
