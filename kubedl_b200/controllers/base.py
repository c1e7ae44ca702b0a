"""Shared body of the four workload reconcilers.

Each reference controller package repeats the same adapter code around the engine
(controllers/pytorch/{pytorchjob_controller,job,pod,service,status}.go and the three
siblings): `Reconcile` (get -> expectations -> default -> ReconcileJobs), the three watches with
their predicates, pod/service claiming through controller-ref managers, status writes, and the
owner-create hook that stamps the Created condition.  That common shape lives here once; the
kind-specific parts (cluster spec, reconcile order, master role, status rules) are the
subclasses in pytorch.py / tensorflow.py / xgboost.py / xdl.py.
"""
from __future__ import annotations

from typing import Dict, List, Optional

from ..api import common as c
from ..api.kinds import SCHEME, KindInfo
from ..apiserver.store import Conflict, NotFound
from ..engine.control import EVENT_NORMAL, ControllerRefManager
from ..engine.interface import ControllerInterface
from ..engine.job_controller import (FAILED_DELETE_JOB_REASON, SUCCESSFUL_DELETE_JOB_REASON,
                                     JobController, JobControllerConfiguration,
                                     recheck_deletion_timestamp)
from ..plugins import gang_schedule
from ..plugins.metrics import JobMetrics
from ..runtime.manager import (Manager, Predicates, Request, Result, enqueue_request_for_object,
                               enqueue_request_for_owner)
from ..utils import status as st
from ..utils.logger import logger


class BaseJobReconciler(ControllerInterface):
    KIND: str = ""
    CONTROLLER_NAME: str = ""

    def __init__(self, mgr: Manager, config: JobControllerConfiguration, metrics_registry=None):
        self.mgr = mgr
        self.client = mgr.client
        self.info: KindInfo = SCHEME.info(self.KIND)
        self.recorder = mgr.get_event_recorder_for(self.CONTROLLER_NAME)
        self.log = logger(f"kubedl.{self.CONTROLLER_NAME}")
        self.metrics = JobMetrics(self.KIND, self.client, registry=metrics_registry)
        gang = None
        if config.enable_gang_scheduling:
            gang = gang_schedule.get(config.gang_scheduler_name)
            if gang is None:
                raise ValueError(f"gang scheduler {config.gang_scheduler_name!r} is not registered "
                                 f"(known: {gang_schedule.REGISTRY.names()})")
        self.ctrl = JobController(self, config, self.client, self.recorder, metrics=self.metrics,
                                  gang_scheduler=gang)
        self.config = config

    # ------------------------------------------------------------------ wiring
    def setup_with_manager(self, mgr: Manager):
        """Three watches: the kind itself, and owned Pods / Services with the expectation
        predicates (reference controllers/pytorch/pytorchjob_controller.go:122-158)."""
        ctl = mgr.new_controller(self.CONTROLLER_NAME, self.reconcile,
                                 max_concurrent_reconciles=self.config.max_concurrent_reconciles)
        ctl.watch(self.KIND, enqueue_request_for_object, Predicates(create=self.on_owner_create))
        owner = enqueue_request_for_owner(self.KIND, is_controller=True)
        ctl.watch("Pod", owner, Predicates(create=self.ctrl.on_pod_create, update=self.ctrl.on_pod_update,
                                           delete=self.ctrl.on_pod_delete))
        ctl.watch("Service", owner, Predicates(create=self.ctrl.on_service_create,
                                               update=self.ctrl.on_service_update,
                                               delete=self.ctrl.on_service_delete))
        self.controller = ctl
        return ctl

    def on_owner_create(self, job: dict) -> bool:
        """Default the new job, stamp the Created condition, count it (reference
        controllers/pytorch/status.go:127-146).  The reference mutates the informer-cache copy and
        relies on the first status write to persist it; we persist it right away."""
        job = c.deepcopy(job)        # the event object is a snapshot shared with the other watchers: never written to
        SCHEME.default(job)
        status = c.status_of(job)
        if not st.has_condition(status, c.JOB_CREATED):
            msg = f"{self.KIND} {c.name_of(job)} is created."
            st.update_job_conditions(status, c.JOB_CREATED, st.JOB_CREATED_REASON, msg)
            # runs on the watch thread, possibly next to the first reconcile: compare-and-swap against
            # what the API server holds now, and add the condition to THAT status instead of replacing it
            for _ in range(4):
                try:
                    fresh = self.client.reader().get(self.KIND, c.namespace_of(job), c.name_of(job))
                except NotFound:
                    return False
                if st.has_condition(c.status_of(fresh), c.JOB_CREATED) or fresh["metadata"].get("deletionTimestamp"):
                    break
                st.update_job_conditions(c.status_of(fresh), c.JOB_CREATED, st.JOB_CREATED_REASON, msg)
                try:
                    self.client.update_status(fresh)
                    break
                except Conflict:
                    continue
                except NotFound:
                    return False
            self.metrics.created_inc()
        return True

    # ------------------------------------------------------------------ Reconcile
    def reconcile(self, req: Request) -> Result:
        try:
            job = self.client.get(self.KIND, req.namespace, req.name)
        except NotFound:
            self.metrics.deleted_inc()
            # the job is gone: so are its pending creations / deletions, and its restart back-off state
            self.ctrl.expectations.delete_for_job(f"{req.namespace}/{req.name}")
            self.ctrl.backoff_states_queue.forget(f"{req.namespace}/{req.name}")
            return Result()
        specs = self.info.replica_specs(job)
        if not self.ctrl.satisfy_expectations(job, specs) or job["metadata"].get("deletionTimestamp"):
            return Result()
        SCHEME.default(job)
        specs = self.info.replica_specs(job)
        try:
            return self.ctrl.reconcile_jobs(job, specs, c.status_of(job), c.run_policy_of(job))
        except NotFound:
            # the job vanished while we were working on it (deleted between the cached read and a
            # write): nothing left to reconcile, the delete event will account for it
            try:
                self.client.reader().get(self.KIND, req.namespace, req.name)
            except NotFound:
                return Result()
            raise

    # ------------------------------------------------------------------ identity
    def controller_name(self) -> str: return self.CONTROLLER_NAME
    def get_api_group_version_kind(self): return (self.info.group, self.info.version, self.info.kind)
    def get_api_group_version(self): return (self.info.group, self.info.version)
    def get_group_name_label_value(self) -> str: return self.info.group_name_label
    def get_default_container_name(self) -> str: return self.info.default_container_name
    def get_default_container_port_name(self) -> str: return self.info.default_port_name
    def get_default_container_port_number(self) -> int: return self.info.default_port

    # ------------------------------------------------------------------ job access
    def get_job_from_informer_cache(self, namespace: str, name: str) -> dict:
        return self.client.get(self.KIND, namespace, name)

    def get_job_from_api_client(self, namespace: str, name: str) -> dict:
        return self.client.reader().get(self.KIND, namespace, name)

    def delete_job(self, job: dict) -> None:
        try:
            self.client.delete(self.KIND, c.namespace_of(job), c.name_of(job))
        except NotFound:
            return
        except Exception as e:
            self.recorder.eventf(job, "Warning", FAILED_DELETE_JOB_REASON, "Error deleting: %s", e)
            raise
        self.recorder.eventf(job, EVENT_NORMAL, SUCCESSFUL_DELETE_JOB_REASON, "Deleted job: %s", c.name_of(job))

    def update_job_status_in_api_server(self, job: dict, job_status: dict) -> None:
        """Status().Update with optimistic concurrency (reference controllers/pytorch/job.go:95-105): the
        write carries the resourceVersion of the object the status was COMPUTED from.  If the job has moved
        on since (a second reconcile worker, another manager, the Created stamp of the watch thread), the
        API server answers Conflict, the reconcile ends there and the request is requeued — a status derived
        from a stale read never overwrites a newer one."""
        obj = c.deepcopy(job)
        obj["status"] = c.deepcopy(job_status)
        written = self.client.update_status(obj)
        rv = (written or {}).get("metadata", {}).get("resourceVersion")
        if rv:      # a later write in the same pass continues from what was just written
            job["metadata"]["resourceVersion"] = rv

    # ------------------------------------------------------------------ children
    def _selector(self, job: dict) -> Dict[str, str]:
        return self.ctrl.gen_labels(c.name_of(job))

    def _can_adopt(self, job: dict):
        def fresh():
            f = self.get_job_from_api_client(c.namespace_of(job), c.name_of(job))
            if c.uid_of(f) != c.uid_of(job):
                raise RuntimeError(f"original {self.KIND} {c.key_of(job)} is gone: got uid {c.uid_of(f)}, "
                                   f"wanted {c.uid_of(job)}")
            return f
        return recheck_deletion_timestamp(fresh)

    def get_pods_for_job(self, job: dict) -> List[dict]:
        """List by {group-name, job-name}, then adopt / release through the ref manager
        (reference controllers/pytorch/pod.go:37-67)."""
        sel = self._selector(job)
        pods = self.client.list("Pod", c.namespace_of(job), sel)
        mgr = ControllerRefManager(self.ctrl.pod_control, job, sel, (self.info.api_version, self.KIND),
                                   self._can_adopt(job), "Pod")
        return mgr.claim(pods)

    def get_services_for_job(self, job: dict) -> List[dict]:
        sel = self._selector(job)
        svcs = self.client.list("Service", c.namespace_of(job), sel)
        mgr = ControllerRefManager(self.ctrl.service_control, job, sel, (self.info.api_version, self.KIND),
                                   self._can_adopt(job), "Service")
        return mgr.claim(svcs)

    def create_pod(self, job: dict, pod: dict) -> None:
        self.client.create(pod)

    def delete_pod(self, job: dict, pod: dict) -> None:
        self.ctrl.pod_control.delete_pod(c.namespace_of(pod), c.name_of(pod), job)

    def create_service(self, job: dict, service: dict) -> None:
        self.client.create(service)

    def delete_service(self, job: dict, name: str, namespace: str) -> None:
        self.ctrl.service_control.delete_service(namespace, name, job)

    def needs_service_for(self, rtype: str) -> bool:
        return True

    # ------------------------------------------------------------------ shared status pieces
    def _mark_failed_or_restarting(self, job: dict, job_status: dict, rtype: str, failed: int, restart: bool,
                                   previous_restarting: bool, previous_failed: bool,
                                   restarting_reason: str = st.JOB_RESTARTING_REASON,
                                   failed_reason: str = st.JOB_FAILED_REASON) -> None:
        name = c.name_of(job)
        if restart:
            msg = f"{self.KIND} {name} is restarting because {failed} {rtype} replica(s) failed."
            self.recorder.event(job, "Warning", restarting_reason, msg)
            st.update_job_conditions(job_status, c.JOB_RESTARTING, st.JOB_RESTARTING_REASON, msg)
            if not previous_restarting:
                self.metrics.failure_inc()
                self.metrics.restart_inc()
        else:
            msg = f"{self.KIND} {name} is failed because {failed} {rtype} replica(s) failed."
            self.recorder.event(job, EVENT_NORMAL, failed_reason, msg)
            if not job_status.get("completionTime"):
                job_status["completionTime"] = c.now_rfc3339()
            st.update_job_conditions(job_status, c.JOB_FAILED, st.JOB_FAILED_REASON, msg)
            if not previous_failed:
                self.metrics.failure_inc()

    def _mark_succeeded(self, job: dict, job_status: dict, msg: str, emit_event: bool = True) -> None:
        if emit_event:
            self.recorder.event(job, EVENT_NORMAL, st.JOB_SUCCEEDED_REASON, msg)
        if not job_status.get("completionTime"):
            job_status["completionTime"] = c.now_rfc3339()
        already = st.is_succeeded(job_status)
        st.update_job_conditions(job_status, c.JOB_SUCCEEDED, st.JOB_SUCCEEDED_REASON, msg)
        if not already:
            self.metrics.success_inc()
