"""The shared job engine: what every workload controller delegates to.

`JobController.reconcile_jobs` is the algorithm of reference pkg/job_controller/job.go:56-266:
gang create -> code-sync injection -> list pods/services -> backoff / active-deadline checks ->
terminal clean-up (CleanPodPolicy, TTL, gang delete, Failed condition) OR per-replica-type
pod/service diff in `get_reconcile_orders()` order -> per-kind status rules -> launch-delay
metrics -> status write when changed.  Pods: pod.go:212-398; services: service.go:188-295;
predicates that feed the expectations: pod.go:53-163, service.go:39-137; naming:
util.go:29-32, job_controller.go:114-136.
"""
from __future__ import annotations

import time
from dataclasses import dataclass
from typing import Dict, List, Optional

from ..api import common as c
from ..apiserver.store import is_already_exists, is_timeout
from ..runtime.expectations import (ControllerExpectations, gen_expectation_pods_key,
                                    gen_expectation_services_key)
from ..runtime.manager import Result
from ..runtime.workqueue import RateLimitingQueue
from ..utils import k8sutil
from ..utils import status as st
from ..utils.logger import logger_for_job, logger_for_replica
from ..utils.train import is_retryable_exit_code
from .control import EVENT_NORMAL, EVENT_WARNING, PodControl, ServiceControl
from .interface import ControllerInterface

FAILED_DELETE_JOB_REASON = "FailedDeleteJob"
SUCCESSFUL_DELETE_JOB_REASON = "SuccessfulDeleteJob"
EXITED_WITH_CODE_REASON = "ExitedWithCode"
POD_TEMPLATE_RESTART_POLICY_REASON = "SettedPodTemplateRestartPolicy"
POD_TEMPLATE_SCHEDULER_NAME_REASON = "SettedPodTemplateSchedulerName"


@dataclass
class JobControllerConfiguration:
    """reference pkg/job_controller/job_controller.go:28-46"""
    enable_gang_scheduling: bool = False
    max_concurrent_reconciles: int = 1
    reconciler_sync_loop_period: float = 15.0   # declared but unused upstream too
    gang_scheduler_name: str = ""
    # pods of one replica type created concurrently, in slow-start batches (1, 2, 4, ...) as the
    # Kubernetes Job controller does; 1 = one at a time in index order like the reference
    # (pod.go:235-281), which costs one API round trip per replica before the last pod exists
    pod_create_parallelism: int = 1


def gen_general_name(job_name: str, rtype: str, index: str) -> str:
    """`<job>-<rtype>-<idx>` with `/` -> `-` (reference util.go:29-32)."""
    return f"{job_name}-{rtype}-{index}".replace("/", "-")


def get_port_from_job(specs: Dict[str, dict], rtype: str, container_name: str, port_name: str) -> int:
    """reference util.go:60-73; raises LookupError when the named port is absent."""
    spec = specs.get(rtype)
    if spec is not None:
        for ctr in (spec.get("template", {}).get("spec", {}).get("containers") or []):
            if ctr.get("name") == container_name:
                for p in ctr.get("ports") or []:
                    if p.get("name") == port_name:
                        return int(p["containerPort"])
    raise LookupError("failed to find the port")


def recheck_deletion_timestamp(get_object):
    """Factory for the adopt guard: re-read the owner uncached and refuse when it is being
    deleted (reference util.go:38-49)."""
    def check():
        obj = get_object()
        if obj.get("metadata", {}).get("deletionTimestamp"):
            raise RuntimeError(f"{c.namespace_of(obj)}/{c.name_of(obj)} has just been deleted at "
                               f"{obj['metadata']['deletionTimestamp']}")
    return check


class _NoMetrics:
    def __getattr__(self, name):
        return lambda *a, **k: None


class JobController:
    def __init__(self, controller: ControllerInterface, config: JobControllerConfiguration, client, recorder,
                 metrics=None, gang_scheduler=None, code_sync=None):
        self.controller = controller
        self.config = config
        self.client = client
        self.recorder = recorder
        self.metrics = metrics if metrics is not None else _NoMetrics()
        self.gang_scheduler = gang_scheduler
        self.expectations = ControllerExpectations()
        self.backoff_states_queue = RateLimitingQueue("backoff-states")
        self.pod_control = PodControl(client, recorder)
        self.service_control = ServiceControl(client, recorder)
        if code_sync is None:
            from ..plugins.code_sync import inject_code_sync_init_containers as code_sync
        self.code_sync = code_sync

    # ------------------------------------------------------------------ identity helpers
    def gen_owner_reference(self, job: dict) -> dict:
        g, v, kind = self.controller.get_api_group_version_kind()
        return {"apiVersion": f"{g}/{v}", "kind": kind, "name": c.name_of(job), "uid": c.uid_of(job),
                "blockOwnerDeletion": True, "controller": True}

    def gen_labels(self, job_name: str) -> Dict[str, str]:
        return {c.GROUP_NAME_LABEL: self.controller.get_group_name_label_value(),
                c.JOB_NAME_LABEL: job_name.replace("/", "-")}

    def resolve_controller_ref(self, namespace: str, ref: Optional[dict]) -> Optional[dict]:
        """Owner lookup by name + UID check (reference job_controller.go:177-193)."""
        if ref is None or ref.get("kind") != self.controller.get_api_group_version_kind()[2]:
            return None
        try:
            job = self.controller.get_job_from_informer_cache(namespace, ref.get("name", ""))
        except Exception:
            return None
        if c.uid_of(job) != ref.get("uid"):
            return None
        return job

    # ------------------------------------------------------------------ gang
    def create_gang(self, job: dict, replicas: Dict[str, dict]):
        gs = self.gang_scheduler
        try:
            return gs.get_gang(c.namespace_of(job), c.name_of(job))
        except Exception:
            pass
        return gs.create_gang(job, replicas)

    def delete_gang(self, job: dict) -> None:
        self.gang_scheduler.delete_gang(c.namespace_of(job), c.name_of(job))

    # ------------------------------------------------------------------ expectations
    def satisfy_expectations(self, job: dict, specs: Dict[str, dict]) -> bool:
        """AND over the pod and service expectation of every replica type
        (reference expectations.go:11-27)."""
        key = c.key_of(job)
        satisfied = True
        for rtype in specs:
            satisfied = satisfied and self.expectations.satisfied_expectations(gen_expectation_pods_key(key, rtype))
            satisfied = satisfied and self.expectations.satisfied_expectations(gen_expectation_services_key(key, rtype))
        return satisfied

    # ------------------------------------------------------------------ event predicates
    def _observe(self, obj: dict, kind_key, creation: bool) -> bool:
        ref = c.controller_ref(obj)
        if ref is None:
            return False
        job = self.resolve_controller_ref(c.namespace_of(obj), ref)
        if job is None:
            return False
        rtype = c.labels_of(obj).get(c.REPLICA_TYPE_LABEL)
        if rtype is None:
            return False
        key = kind_key(c.key_of(job), rtype)
        if creation:
            self.expectations.creation_observed(key)
        else:
            self.expectations.deletion_observed(key)
        return True

    def on_pod_create(self, pod: dict) -> bool:
        if pod.get("metadata", {}).get("deletionTimestamp"):
            return False
        return self._observe(pod, gen_expectation_pods_key, True)

    def on_pod_update(self, old: dict, new: dict) -> bool:
        if old.get("metadata", {}).get("resourceVersion") == new.get("metadata", {}).get("resourceVersion"):
            return False
        ref = c.controller_ref(new)
        return ref is not None and self.resolve_controller_ref(c.namespace_of(new), ref) is not None

    def on_pod_delete(self, pod: dict) -> bool:
        return self._observe(pod, gen_expectation_pods_key, False)

    def on_service_create(self, svc: dict) -> bool:
        return self._observe(svc, gen_expectation_services_key, True)

    def on_service_update(self, old: dict, new: dict) -> bool:
        if old.get("metadata", {}).get("resourceVersion") == new.get("metadata", {}).get("resourceVersion"):
            return False
        ref = c.controller_ref(new)
        return ref is not None and self.resolve_controller_ref(c.namespace_of(new), ref) is not None

    def on_service_delete(self, svc: dict) -> bool:
        return self._observe(svc, gen_expectation_services_key, False)

    # ------------------------------------------------------------------ run policies
    def delete_pods_and_services(self, run_policy: dict, job: dict, pods: List[dict]) -> None:
        """reference job.go:29-52"""
        if not pods:
            return
        policy = run_policy.get("cleanPodPolicy") or c.CLEAN_POD_POLICY_NONE
        if policy == c.CLEAN_POD_POLICY_NONE:
            return
        for pod in pods:
            if policy == c.CLEAN_POD_POLICY_RUNNING and k8sutil.phase_of(pod) != c.POD_RUNNING:
                continue
            self.controller.delete_pod(job, pod)
            # pod and service share a name
            self.controller.delete_service(job, c.name_of(pod), c.namespace_of(pod))

    def past_active_deadline(self, run_policy: dict, job_status: dict) -> bool:
        """reference job.go:269-278"""
        ads = run_policy.get("activeDeadlineSeconds")
        start = c.parse_time(job_status.get("startTime"))
        if ads is None or start is None:
            return False
        return time.time() - start >= float(ads)

    def past_backoff_limit(self, job_name: str, run_policy: dict, replicas: Dict[str, dict],
                           pods: List[dict]) -> bool:
        """Σ restartCounts of Running pods of OnFailure/Always replica types vs the limit
        (reference job.go:282-319)."""
        limit = run_policy.get("backoffLimit")
        if limit is None:
            return False
        result = 0
        for rtype, spec in replicas.items():
            if spec.get("restartPolicy") not in (c.RESTART_POLICY_ON_FAILURE, c.RESTART_POLICY_ALWAYS):
                continue
            for po in self.filter_pods_for_replica_type(pods, rtype.lower()):
                if k8sutil.phase_of(po) != c.POD_RUNNING:
                    continue
                stt = po.get("status") or {}
                for cs in (stt.get("initContainerStatuses") or []) + (stt.get("containerStatuses") or []):
                    result += int(cs.get("restartCount", 0))
        if int(limit) == 0:
            return result > 0
        return result >= int(limit)

    def cleanup_job(self, run_policy: dict, job_status: dict, job: dict) -> Result:
        """TTL after finished (reference job.go:321-345)."""
        ttl = run_policy.get("ttlSecondsAfterFinished")
        if ttl is None:
            return Result()
        done = c.parse_time(job_status.get("completionTime"))
        if done is None:
            raise RuntimeError(f"cleanup Job {c.name_of(job)}, but job has CompletionTime not set")
        delete_at = done + float(ttl)
        now = time.time()
        if now > delete_at or float(ttl) == 0:
            self.controller.delete_job(job)
            return Result()
        return Result(requeue=True, requeue_after=delete_at - now)

    # ------------------------------------------------------------------ the main algorithm
    def reconcile_jobs(self, job: dict, replicas: Dict[str, dict], job_status: dict, run_policy: dict) -> Result:
        job_key = c.key_of(job)
        job_name = c.name_of(job)
        result = Result()
        try:
            result = self._reconcile_jobs(job, job_key, job_name, replicas, job_status, run_policy)
        except Exception:
            self.backoff_states_queue.add_rate_limited(job_key)
            raise
        if result.requeue:
            self.backoff_states_queue.add_rate_limited(job_key)
        else:
            self.backoff_states_queue.forget(job_key)
        return result

    def _reconcile_jobs(self, job, job_key, job_name, replicas, job_status, run_policy) -> Result:
        result = Result()
        if self.config.enable_gang_scheduling and self.gang_scheduler is not None:
            self.create_gang(job, replicas)

        old_status = c.deepcopy(job_status)
        self.code_sync(job, replicas)

        pods = self.controller.get_pods_for_job(job)
        services = self.controller.get_services_for_job(job)

        previous_retry = self.backoff_states_queue.num_requeues(job_key)
        active_pods = k8sutil.filter_active_pods(pods)
        active = len(active_pods)
        failed = k8sutil.filter_pod_count(pods, c.POD_FAILED)
        total = k8sutil.get_total_replicas(replicas)
        prev_failed = k8sutil.get_total_failed_replicas(job_status.get("replicaStatuses"))

        failure_message = ""
        job_exceeds_limit = False
        exceeds_backoff = False
        past_backoff = False
        if run_policy.get("backoffLimit") is not None:
            has_new_failure = failed > prev_failed
            exceeds_backoff = has_new_failure and active != total and \
                (previous_retry + 1 > int(run_policy["backoffLimit"]))
            past_backoff = self.past_backoff_limit(job_name, run_policy, replicas, pods)

        if exceeds_backoff or past_backoff:
            job_exceeds_limit = True
            failure_message = f"Job {job_name} has failed because it has reached the specified backoff limit"
        elif self.past_active_deadline(run_policy, job_status):
            failure_message = f"Job {job_name} has failed because it was active longer than specified deadline"
            job_exceeds_limit = True
            job_status["completionTime"] = c.now_rfc3339()

        if st.is_succeeded(job_status) or st.is_failed(job_status) or job_exceeds_limit:
            self.delete_pods_and_services(run_policy, job, pods)
            result = self.cleanup_job(run_policy, job_status, job)
            if self.config.enable_gang_scheduling and self.gang_scheduler is not None:
                self.recorder.event(job, EVENT_NORMAL, "JobTerminated", "Job has been terminated. Deleting PodGroup")
                try:
                    self.delete_gang(job)
                except Exception as e:
                    self.recorder.eventf(job, EVENT_WARNING, "FailedDeletePodGroup", "Error deleting: %s", e)
                    raise
                self.recorder.eventf(job, EVENT_NORMAL, "SuccessfulDeletePodGroup", "Deleted PodGroup: %s", job_name)
            if job_exceeds_limit:
                self.recorder.event(job, EVENT_NORMAL, st.JOB_FAILED_REASON, failure_message)
                if not job_status.get("completionTime"):
                    job_status["completionTime"] = c.now_rfc3339()
                st.update_job_conditions(job_status, c.JOB_FAILED, st.JOB_FAILED_REASON, failure_message)
            if st.is_succeeded(job_status):
                # pods may be gone by now: fold whatever was still active into succeeded
                for rs in (job_status.get("replicaStatuses") or {}).values():
                    rs["succeeded"] = int(rs.get("succeeded", 0)) + int(rs.get("active", 0))
                    rs["active"] = 0
            if old_status != job_status:
                self.controller.update_job_status_in_api_server(job, job_status)
            return result

        restart = [False]
        for rtype in self.controller.get_reconcile_orders():
            spec = replicas.get(rtype)
            if spec is None:
                continue
            self.reconcile_pods(job, job_status, pods, rtype, spec, replicas, restart)
            # the reference hard-codes "PyTorch needs a Service only for its Master" in the
            # generic engine (job.go:223-227); here the controller answers the question
            if not self.needs_service(rtype):
                continue
            self.reconcile_services(job, services, rtype, spec)

        self.controller.update_job_status(job, replicas, job_status, restart[0])

        # reference condition (job.go:243-245); the metric itself observes once per job UID
        if st.is_created(old_status) and st.is_running(job_status):
            self.metrics.first_pod_launch_delay_seconds(active_pods, job, job_status)
        if k8sutil.get_total_active_replicas(job_status.get("replicaStatuses")) == total and \
                k8sutil.get_total_active_replicas(old_status.get("replicaStatuses")) < total and \
                not st.is_restarting(old_status):
            self.metrics.all_pods_launch_delay_seconds(pods, job, job_status)

        if old_status != job_status:
            self.controller.update_job_status_in_api_server(job, job_status)
        # The reference has no timer for activeDeadlineSeconds: the deadline is noticed only when some
        # other event triggers a reconcile (job.go:151-156; the only RequeueAfter is the TTL one).
        # Fixed here: come back when the deadline expires.
        ads = run_policy.get("activeDeadlineSeconds")
        start = c.parse_time(job_status.get("startTime"))
        if ads is not None and start is not None and result.requeue_after <= 0:
            # requeue=False: a timer wake-up is not a retry and must not count towards backoffLimit
            result = Result(requeue=False, requeue_after=max(start + float(ads) - time.time(), 0.0) + 0.05)
        return result

    def needs_service(self, rtype: str) -> bool:
        fn = getattr(self.controller, "needs_service_for", None)
        return True if fn is None else bool(fn(rtype))

    # ------------------------------------------------------------------ pods
    @staticmethod
    def filter_pods_for_replica_type(pods: List[dict], rtype_lower: str) -> List[dict]:
        return [p for p in pods if c.labels_of(p).get(c.REPLICA_TYPE_LABEL) == rtype_lower]

    filter_services_for_replica_type = filter_pods_for_replica_type

    @staticmethod
    def get_pod_slices(objs: List[dict], replicas: int, log=None) -> List[List[dict]]:
        """Bucket by the replica-index label; out-of-range indices are only logged — there is no
        scale-down (reference pod.go:189-208)."""
        slices: List[List[dict]] = [[] for _ in range(replicas)]
        for o in objs:
            raw = c.labels_of(o).get(c.REPLICA_INDEX_LABEL)
            if raw is None:
                if log:
                    log.warning("The pod do not have the index label.")
                continue
            try:
                idx = int(raw)
            except ValueError:
                continue
            if idx < 0 or idx >= replicas:
                if log:
                    log.warning("The label index is not expected: %d", idx)
            else:
                slices[idx].append(o)
        return slices

    get_service_slices = get_pod_slices

    def reconcile_pods(self, job: dict, job_status: dict, pods: List[dict], rtype: str, spec: dict,
                       replicas: Dict[str, dict], restart: List[bool]) -> None:
        rt = rtype.lower()
        log = logger_for_replica(job, rt)
        pods = self.filter_pods_for_replica_type(pods, rt)
        num = c.replicas_of(spec)
        initialize_replica_statuses(job_status, rtype)
        slices = self.get_pod_slices(pods, num, log)

        def create(index: int) -> None:
            master = self.controller.is_master_role(replicas, rtype, index)
            try:
                self.create_new_pod(job, rt, str(index), spec, master, replicas)
            except Exception as e:
                if is_already_exists(e):
                    # the cache missed an earlier create: observe it ourselves so the queue
                    # does not stall on the expectation (reference pod.go:265-278)
                    # (create_new_pod has taken the failed create out of the pod count already)
                    key = c.key_of(job)
                    self.expectations.creation_observed(gen_expectation_services_key(key, rt))
                raise

        missing = [i for i, sl in enumerate(slices) if len(sl) == 0]
        parallel = max(1, int(getattr(self.config, "pod_create_parallelism", 1) or 1))
        batched = parallel > 1 and len(missing) > 1
        if batched:
            self._create_in_batches(missing, create, parallel)
        for index, pod_slice in enumerate(slices):
            if len(pod_slice) > 1:
                log.warning("We have too many pods for %s %d", rt, index)
            elif len(pod_slice) == 0:
                if not batched:
                    create(index)
            else:
                pod = pod_slice[0]
                exit_code = 0xBEEF
                for cs in (pod.get("status") or {}).get("containerStatuses") or []:
                    term = (cs.get("state") or {}).get("terminated")
                    if cs.get("name") == self.controller.get_default_container_name() and term is not None:
                        exit_code = int(term.get("exitCode", 0))
                        self.recorder.eventf(job, EVENT_NORMAL, EXITED_WITH_CODE_REASON,
                                             "Pod: %s.%s exited with code %s", c.namespace_of(pod),
                                             c.name_of(pod), exit_code)
                        break
                if spec.get("restartPolicy") == c.RESTART_POLICY_EXIT_CODE:
                    if k8sutil.phase_of(pod) == c.POD_FAILED and is_retryable_exit_code(exit_code):
                        self.controller.delete_pod(job, pod)
                        restart[0] = True
                update_job_replica_statuses(job_status, rtype, pod)

    @staticmethod
    def _create_in_batches(indices: List[int], create, parallel: int) -> None:
        """Slow start: 1, 2, 4, ... creates at a time (capped at `parallel`); a failing batch stops
        the ramp — when the API server rejects pods, the rest would be rejected too — and the
        first error is raised once the batch has finished."""
        from concurrent.futures import ThreadPoolExecutor
        todo = list(indices)
        size = 1
        with ThreadPoolExecutor(max_workers=parallel, thread_name_prefix="pod-create") as pool:
            while todo:
                batch, todo = todo[:size], todo[size:]
                errors = [f.exception() for f in [pool.submit(create, i) for i in batch]]
                errors = [e for e in errors if e is not None]
                if errors:
                    raise errors[0]
                size = min(size * 2, parallel)

    def create_new_pod(self, job: dict, rt: str, index: str, spec: dict, master_role: bool,
                       replicas: Dict[str, dict]) -> None:
        job_key = c.key_of(job)
        exp_key = gen_expectation_pods_key(job_key, rt)
        self.expectations.expect_one_more_creation(exp_key)
        try:
            labels = self.gen_labels(c.name_of(job))
            labels[c.REPLICA_TYPE_LABEL] = rt
            labels[c.REPLICA_INDEX_LABEL] = index
            if master_role:
                labels[c.JOB_ROLE_LABEL] = "master"
            template = c.deepcopy(spec.get("template") or {})
            tm = template.setdefault("metadata", {})
            tm["name"] = gen_general_name(c.name_of(job), rt, index)
            tm.setdefault("labels", {})
            if tm["labels"] is None:
                tm["labels"] = {}
            tm["labels"].update(labels)
            template.setdefault("spec", {})
            self.controller.set_cluster_spec(job, template, rt, index)
            if template["spec"].get("restartPolicy"):
                msg = "Restart policy in pod template will be overwritten by restart policy in replica spec"
                self.recorder.event(job, EVENT_WARNING, POD_TEMPLATE_RESTART_POLICY_REASON, msg)
            set_restart_policy(template, spec)
            if self.config.enable_gang_scheduling and self.gang_scheduler is not None:
                gang = self.gang_scheduler.get_gang(c.namespace_of(job), c.name_of(job))
                self.gang_scheduler.bind_pod_to_gang(template, gang)
        except Exception:
            self.expectations.creation_observed(exp_key)   # nothing was sent: nothing to wait for
            raise
        ref = self.gen_owner_reference(job)
        try:
            pod = self.pod_control_template_to_pod(template, job, ref)
            self.controller.create_pod(job, pod)
        except Exception as e:
            if is_timeout(e):
                return  # the create may still land; the expectation/TTL path resolves it
            # no ADDED event will come for this one (AlreadyExists included: that pod's event was
            # either seen already or belongs to an earlier expectation)
            self.expectations.creation_observed(exp_key)
            self.recorder.eventf(job, EVENT_WARNING, "FailedCreatePod", "Error creating: %s", e)
            raise
        self.recorder.eventf(job, EVENT_NORMAL, "SuccessfulCreatePod", "Created pod: %s", tm["name"])

    @staticmethod
    def pod_control_template_to_pod(template: dict, job: dict, ref: dict) -> dict:
        from .control import get_pod_from_template, validate_controller_ref
        validate_controller_ref(ref)
        pod = get_pod_from_template(template, job, ref)
        pod["metadata"]["namespace"] = c.namespace_of(job)
        if not pod["metadata"].get("labels"):
            raise ValueError("unable to create pods, no labels")
        return pod

    # ------------------------------------------------------------------ services
    def reconcile_services(self, job: dict, services: List[dict], rtype: str, spec: dict) -> None:
        rt = rtype.lower()
        log = logger_for_replica(job, rt)
        services = self.filter_services_for_replica_type(services, rt)
        for index, sl in enumerate(self.get_service_slices(services, c.replicas_of(spec), log)):
            if len(sl) > 1:
                log.warning("We have too many services for %s %d", rt, index)
            elif len(sl) == 0:
                self.create_new_service(job, rtype, spec, str(index))

    def get_port_from_spec(self, spec: dict) -> int:
        """reference service.go:221-234"""
        for ctr in (spec.get("template", {}).get("spec", {}).get("containers") or []):
            if ctr.get("name") == self.controller.get_default_container_name():
                for p in ctr.get("ports") or []:
                    if p.get("name") == self.controller.get_default_container_port_name():
                        return int(p["containerPort"])
        raise LookupError("failed to find the port")

    def create_new_service(self, job: dict, rtype: str, spec: dict, index: str) -> None:
        rt = rtype.lower()
        exp_key = gen_expectation_services_key(c.key_of(job), rt)
        self.expectations.expect_one_more_creation(exp_key)
        labels = self.gen_labels(c.name_of(job))
        labels[c.REPLICA_TYPE_LABEL] = rt
        labels[c.REPLICA_INDEX_LABEL] = index
        try:
            port = self.get_port_from_spec(spec)
        except LookupError:
            self.expectations.creation_observed(exp_key)
            raise
        name = gen_general_name(c.name_of(job), rt, index)
        svc = {"apiVersion": "v1", "kind": "Service",
               "metadata": {"name": name, "namespace": c.namespace_of(job), "labels": dict(labels),
                            "ownerReferences": [self.gen_owner_reference(job)]},
               "spec": {"clusterIP": "None", "selector": dict(labels),
                        "ports": [{"name": self.controller.get_default_container_port_name(), "port": port}]}}
        try:
            self.controller.create_service(job, svc)
        except Exception as e:
            if is_timeout(e):
                return
            self.expectations.creation_observed(exp_key)
            self.recorder.eventf(job, EVENT_WARNING, "FailedCreateService", "Error creating: %s", e)
            raise RuntimeError(f"unable to create services: {e}") from e
        self.recorder.eventf(job, EVENT_NORMAL, "SuccessfulCreateService", "Created service: %s", name)


# ---------------------------------------------------------------------- status helpers
def initialize_replica_statuses(job_status: dict, rtype: str) -> None:
    """Counters of a type are reset at every reconcile (reference status.go:9-16)."""
    job_status.setdefault("replicaStatuses", {})
    if job_status["replicaStatuses"] is None:
        job_status["replicaStatuses"] = {}
    job_status["replicaStatuses"][rtype] = c.new_replica_status()


def update_job_replica_statuses(job_status: dict, rtype: str, pod: dict) -> None:
    """Running -> active, Succeeded, Failed (reference status.go:18-27)."""
    rs = job_status["replicaStatuses"][rtype]
    phase = k8sutil.phase_of(pod)
    if phase == c.POD_RUNNING:
        rs["active"] += 1
    elif phase == c.POD_SUCCEEDED:
        rs["succeeded"] += 1
    elif phase == c.POD_FAILED:
        rs["failed"] += 1


def set_restart_policy(template: dict, spec: dict) -> None:
    """ExitCode -> Never, everything else verbatim, including the empty policy XGBoost leaves
    (reference pod.go:435-442)."""
    rp = spec.get("restartPolicy") or ""
    template.setdefault("spec", {})["restartPolicy"] = c.RESTART_POLICY_NEVER if rp == c.RESTART_POLICY_EXIT_CODE else rp
