"""Job engine parity tests (reference pkg/job_controller/job_test.go:15-216, pod_test.go:12-64,
status_test.go:11-40, service_ref_manager_test.go:29-200, pod_control_test.go, service_control_test.go)
plus the full-loop reconcile tests the reference never had (SURVEY.md §4 "Not tested at all")."""
import time

import pytest

from cp_helpers import Cluster, tmpl
from kubedl_b200.api import common as c
from kubedl_b200.api import kinds as k
from kubedl_b200.apiserver.store import APIServer, Timeout
from kubedl_b200.engine import control
from kubedl_b200.engine.job_controller import (initialize_replica_statuses, set_restart_policy,
                                               update_job_replica_statuses)
from kubedl_b200.runtime.expectations import ControllerExpectations, gen_expectation_pods_key
from kubedl_b200.runtime.manager import EventRecorder, Manager
from kubedl_b200.runtime.workqueue import RateLimitingQueue


def _tfjob(name="tf", workers=2, ps=0, **spec):
    specs = {"Worker": {"replicas": workers, "restartPolicy": "Never", "template": tmpl("tensorflow")}}
    if ps:
        specs["PS"] = {"replicas": ps, "restartPolicy": "Never", "template": tmpl("tensorflow")}
    return k.new_job("TFJob", name, "ns", specs, **spec)


# ---------------------------------------------------------------- small units
@pytest.mark.parametrize("policy,expected", [("ExitCode", "Never"), ("Never", "Never"), ("Always", "Always"),
                                             ("OnFailure", "OnFailure"), (None, "")])
def test_set_restart_policy(policy, expected):
    t = {"spec": {}}
    set_restart_policy(t, {"restartPolicy": policy} if policy else {})
    assert t["spec"]["restartPolicy"] == expected


def test_replica_status_counting():
    s = c.new_job_status()
    initialize_replica_statuses(s, "Worker")
    for phase in ("Running", "Running", "Succeeded", "Failed", "Pending"):
        update_job_replica_statuses(s, "Worker", {"status": {"phase": phase}})
    assert s["replicaStatuses"]["Worker"] == {"active": 2, "succeeded": 1, "failed": 1}
    initialize_replica_statuses(s, "Worker")   # reset at each reconcile
    assert s["replicaStatuses"]["Worker"] == {"active": 0, "succeeded": 0, "failed": 0}


def test_expectations_semantics():
    t = [0.0]
    e = ControllerExpectations(clock=lambda: t[0], ttl=300)
    key = gen_expectation_pods_key("ns/j", "Worker")
    assert key == "ns/j/worker/pods"
    assert e.satisfied_expectations(key)                 # never set -> sync
    e.expect_creations(key, 1)
    assert not e.satisfied_expectations(key)
    e.expect_creations(key, 1)                           # overwrite, not accumulate (SetExpectations)
    e.creation_observed(key)
    assert e.satisfied_expectations(key)
    e.expect_deletions(key, 2)
    e.deletion_observed(key)
    assert not e.satisfied_expectations(key)
    t[0] = 301.0                                         # expired after 5 min
    assert e.satisfied_expectations(key)


def test_workqueue_dedup_and_backoff():
    t = [0.0]
    q = RateLimitingQueue(clock=lambda: t[0])
    q.add("a"); q.add("a"); q.add("b")
    assert len(q) == 2
    item, _ = q.get(block=False)
    q.add(item)                                          # re-added while processing -> queued after done()
    assert len(q) == 1
    q.done(item)
    assert len(q) == 2
    assert q.num_requeues("x") == 0
    q.add_rate_limited("x"); q.add_rate_limited("x")
    assert q.num_requeues("x") == 2
    assert q.get(block=False)[0] in ("b", "a")
    q.forget("x")
    assert q.num_requeues("x") == 0
    t[0] = 10.0
    names = set()
    while True:
        it, _ = q.get(block=False)
        if it is None:
            break
        names.add(it); q.done(it)
    assert "x" in names


def test_pod_and_service_control_create_with_controller_ref():
    mgr = Manager()
    rec = EventRecorder(mgr.api, "test")
    parent = mgr.api.create(k.new_job("TFJob", "j", "ns", {}))
    ref = {"apiVersion": parent["apiVersion"], "kind": "TFJob", "name": "j", "uid": parent["metadata"]["uid"],
           "controller": True, "blockOwnerDeletion": True}
    pc = control.PodControl(mgr.client, rec)
    template = {"metadata": {"name": "p", "labels": {"a": "b"}, "annotations": {"x": "y"}},
                "spec": {"containers": [{"name": "c", "image": "i"}]}}
    pod = pc.create_pods_with_controller_ref("ns", template, parent, ref)
    got = mgr.api.get("Pod", "ns", "p")
    assert got["metadata"]["labels"] == {"a": "b"} and got["metadata"]["ownerReferences"] == [ref]
    assert got["spec"] == template["spec"]
    with pytest.raises(ValueError):
        pc.create_pods_with_controller_ref("ns", template, parent, {**ref, "controller": False})
    with pytest.raises(ValueError):
        pc.create_pods_with_controller_ref("ns", {"metadata": {"name": "q"}, "spec": {}}, parent, ref)
    sc = control.ServiceControl(mgr.client, rec)
    svc = {"metadata": {"name": "s", "labels": {"a": "b"}}, "spec": {"clusterIP": "None"}}
    sc.create_services_with_controller_ref("ns", svc, parent, ref)
    assert mgr.api.get("Service", "ns", "s")["metadata"]["ownerReferences"][0]["uid"] == ref["uid"]
    sc.delete_service("ns", "s", parent)
    sc.delete_service("ns", "s", parent)   # deleting a missing service is not an error
    assert control.SUCCESSFUL_CREATE_POD_REASON in rec.reasons()
    assert control.SUCCESSFUL_DELETE_SERVICE_REASON in rec.reasons()


def _svc(name, labels, owner=None, deleting=False):
    m = {"name": name, "namespace": "ns", "labels": labels, "uid": "u-" + name}
    if owner:
        m["ownerReferences"] = [{"apiVersion": "kubeflow.org/v1", "kind": "TFJob", "name": owner["metadata"]["name"],
                                 "uid": owner["metadata"]["uid"], "controller": True, "blockOwnerDeletion": True}]
    if deleting:
        m["deletionTimestamp"] = c.now_rfc3339()
    return {"kind": "Service", "metadata": m, "spec": {}}


def test_service_ref_manager_claims():
    """claim by label; a deleting controller cannot adopt; others' services are left alone;
    ours is released when labels stop matching (reference service_ref_manager_test.go:29-200)."""
    job = {"kind": "TFJob", "metadata": {"name": "j", "namespace": "ns", "uid": "J"}}
    other = {"kind": "TFJob", "metadata": {"name": "o", "namespace": "ns", "uid": "O"}}
    sel = {"foo": "bar"}
    fake = control.FakeServiceControl()
    ok = lambda: None
    m = control.ControllerRefManager(fake, job, sel, ("kubeflow.org/v1", "TFJob"), ok, "Service")
    got = m.claim([_svc("s1", sel), _svc("s2", {"foo": "nope"})])
    assert [c.name_of(s) for s in got] == ["s1"] and len(fake.patches) == 1      # adopted s1
    fake.clear()
    deleting = {"kind": "TFJob", "metadata": {"name": "j", "namespace": "ns", "uid": "J",
                                               "deletionTimestamp": c.now_rfc3339()}}
    m = control.ControllerRefManager(fake, deleting, sel, ("kubeflow.org/v1", "TFJob"), ok, "Service")
    assert m.claim([_svc("s1", sel)]) == [] and not fake.patches
    m = control.ControllerRefManager(fake, job, sel, ("kubeflow.org/v1", "TFJob"), ok, "Service")
    assert m.claim([_svc("s1", sel, owner=other)]) == [] and not fake.patches     # owned by someone else
    got = m.claim([_svc("mine", sel, owner=job), _svc("drifted", {"foo": "x"}, owner=job)])
    assert [c.name_of(s) for s in got] == ["mine"] and len(fake.patches) == 1     # released "drifted"
    fake.clear()

    def refuse():
        raise RuntimeError("owner is gone")
    m = control.ControllerRefManager(fake, job, sel, ("kubeflow.org/v1", "TFJob"), refuse, "Service")
    with pytest.raises(RuntimeError):
        m.claim([_svc("s1", sel)])
    assert m.claim([_svc("deleting", sel, deleting=True)]) == []                  # never adopt a dying child


# ---------------------------------------------------------------- run policies (job_test.go)
@pytest.mark.parametrize("policy,left", [("Running", {"tf-worker-1", "tf-worker-2"}), ("All", set()),
                                         ("None", {"tf-worker-0", "tf-worker-1", "tf-worker-2"})])
def test_clean_pod_policy(policy, left):
    cl = Cluster()
    cl.submit(_tfjob(workers=3, cleanPodPolicy=policy))
    cl.kubelet.set_phase("ns", "tf-worker-0", "Running")
    cl.kubelet.set_phase("ns", "tf-worker-1", "Succeeded")
    cl.kubelet.set_phase("ns", "tf-worker-2", "Failed", exit_code=1)     # permanent -> job Failed
    cl.settle()
    assert ("Failed", "True") in cl.conditions("TFJob", "ns", "tf")
    assert {c.name_of(p) for p in cl.pods("ns")} == left
    assert {c.name_of(s) for s in cl.api.list("Service", "ns")} == left


def test_past_backoff_limit_counts_restarts_of_running_pods():
    cl = Cluster()
    r = cl.recs["TFJob"].ctrl
    specs = {"Worker": {"replicas": 2, "restartPolicy": "OnFailure"}, "PS": {"replicas": 1, "restartPolicy": "Never"}}
    pod = lambda rt, rc, phase="Running": {"metadata": {"labels": {c.REPLICA_TYPE_LABEL: rt}},
                                           "status": {"phase": phase, "containerStatuses": [{"restartCount": rc}]}}
    assert not r.past_backoff_limit("j", {"backoffLimit": 0}, specs, [pod("worker", 0), pod("worker", 0)])
    assert r.past_backoff_limit("j", {"backoffLimit": 0}, specs, [pod("worker", 1)])
    assert not r.past_backoff_limit("j", {"backoffLimit": 3}, specs, [pod("worker", 1), pod("worker", 1)])
    assert r.past_backoff_limit("j", {"backoffLimit": 3}, specs, [pod("worker", 2), pod("worker", 1)])
    assert not r.past_backoff_limit("j", {"backoffLimit": 1}, specs, [pod("ps", 5)])          # Never: not counted
    assert not r.past_backoff_limit("j", {"backoffLimit": 1}, specs, [pod("worker", 5, "Failed")])
    assert not r.past_backoff_limit("j", {}, specs, [pod("worker", 5)])


def test_past_active_deadline():
    r = Cluster().recs["TFJob"].ctrl
    now = c.now_rfc3339()
    assert r.past_active_deadline({"activeDeadlineSeconds": 0}, {"startTime": now})
    assert not r.past_active_deadline({"activeDeadlineSeconds": 2}, {"startTime": now})
    assert not r.past_active_deadline({}, {"startTime": now})
    assert not r.past_active_deadline({"activeDeadlineSeconds": 0}, {})


def test_cleanup_job_ttl():
    cl = Cluster()
    cl.submit(_tfjob(workers=1, ttlSecondsAfterFinished=0))
    cl.kubelet.set_phase("ns", "tf-worker-0", "Succeeded")
    cl.settle()
    assert cl.api.list("TFJob", "ns") == []                    # TTL 0 -> deleted right after finishing
    assert cl.pods("ns") == []                                  # owner GC took the pods
    cl2 = Cluster()
    cl2.submit(_tfjob(workers=1, ttlSecondsAfterFinished=3600))
    cl2.kubelet.set_phase("ns", "tf-worker-0", "Succeeded")
    cl2.settle()
    assert len(cl2.api.list("TFJob", "ns")) == 1               # requeued for later, still there
    assert cl2.recs["TFJob"].controller.queue.pending_delayed() >= 1


def test_active_deadline_marks_failed_and_cleans_up():
    cl = Cluster()
    cl.submit(_tfjob(workers=2, activeDeadlineSeconds=0, cleanPodPolicy="All"))
    cl.kubelet.set_all("ns", "Running")
    cl.settle()
    conds = cl.conditions("TFJob", "ns", "tf")
    assert conds[-1] == ("Failed", "True")
    job = cl.job("TFJob", "ns", "tf")
    assert "deadline" in job["status"]["conditions"][-1]["message"] and job["status"]["completionTime"]
    assert cl.pods("ns") == []


def test_active_deadline_fires_without_any_other_event():
    """The reference only notices an expired deadline when something else triggers a reconcile
    (SURVEY.md §2.6 item 15); here the reconcile asks to be woken up when the deadline passes."""
    import time
    cl = Cluster()
    cl.submit(_tfjob(workers=1, activeDeadlineSeconds=2, cleanPodPolicy="All"))   # startTime has 1 s resolution
    cl.kubelet.set_all("ns", "Running")
    cl.settle()
    assert cl.conditions("TFJob", "ns", "tf")[-1] == ("Running", "True")
    assert cl.recs["TFJob"].controller.queue.pending_delayed() >= 1
    time.sleep(2.2)
    cl.settle()                                   # no pod / job event in between
    assert cl.conditions("TFJob", "ns", "tf")[-1] == ("Failed", "True")
    assert cl.pods("ns") == []


def test_backoff_limit_exceeded_by_restart_counts():
    cl = Cluster()
    job = k.new_job("TFJob", "tf", "ns", {"Worker": {"replicas": 1, "restartPolicy": "OnFailure",
                                                      "template": tmpl("tensorflow")}}, backoffLimit=2)
    cl.submit(job)
    cl.kubelet.set_phase("ns", "tf-worker-0", "Running", restart_count=2)
    cl.settle()
    assert cl.conditions("TFJob", "ns", "tf")[-1] == ("Failed", "True")
    assert "backoff limit" in cl.job("TFJob", "ns", "tf")["status"]["conditions"][-1]["message"]


# ---------------------------------------------------------------- engine behaviours
def test_already_exists_recovery_does_not_stall_expectations():
    """A pod the cache did not know about: create fails with AlreadyExists, the engine observes the
    creation itself and the next pass proceeds (reference pod.go:254-279)."""
    cl = Cluster()
    job = cl.api.create(_tfjob(workers=1))
    r = cl.recs["TFJob"]
    from kubedl_b200.apiserver.store import AlreadyExists
    cl.api.fail_next[("create", "Pod")] = AlreadyExists("pod exists")
    cl.settle()
    key = gen_expectation_pods_key("ns/tf", "worker")
    assert r.ctrl.expectations.satisfied_expectations(key)
    cl.mgr.run_until_idle(advance_delayed=True)
    assert [c.name_of(p) for p in cl.pods("ns")] == ["tf-worker-0"]


def test_expectations_count_every_create_in_flight():
    """Eight creates mean eight ADDED events to wait for, not one (the reference's
    `ExpectCreations(key, 1)` per pod overwrites, pod.go:329-333; see docs/parity_notes.md #5)."""
    from kubedl_b200.runtime.expectations import ControllerExpectations
    now = [0.0]
    e = ControllerExpectations(clock=lambda: now[0], ttl=300)
    for _ in range(3):
        e.expect_one_more_creation("k")
    e.creation_observed("k")
    assert not e.satisfied_expectations("k")          # two still unseen
    e.creation_observed("k")
    e.creation_observed("k")
    assert e.satisfied_expectations("k")
    e.expect_one_more_creation("k")                   # a satisfied record starts again at one
    assert e.get("k").add == 1
    now[0] = 301.0
    assert e.satisfied_expectations("k")              # expired: resync
    e.expect_one_more_creation("k")                   # an expired record is not grown
    assert e.get("k").add == 1 and not e.satisfied_expectations("k")


def test_failed_create_does_not_leave_an_expectation_behind():
    cl = Cluster()
    cl.api.create(_tfjob(workers=2))
    cl.api.fail_next[("create", "Pod")] = RuntimeError("admission webhook said no")
    cl.settle()
    r = cl.recs["TFJob"]
    assert r.ctrl.expectations.satisfied_expectations(gen_expectation_pods_key("ns/tf", "worker"))
    cl.mgr.run_until_idle(advance_delayed=True)
    assert [c.name_of(p) for p in cl.pods("ns")] == ["tf-worker-0", "tf-worker-1"]


def test_timeout_on_create_is_not_an_error():
    cl = Cluster()
    cl.api.create(_tfjob(workers=1))
    cl.api.fail_next[("create", "Pod")] = Timeout("slow apiserver")
    cl.settle()
    assert cl.recs["TFJob"].controller.error_count == 0


def test_adoption_of_orphan_pod_and_no_scale_down():
    cl = Cluster()
    cl.submit(_tfjob(workers=2))
    # an orphan with matching labels is adopted on the next reconcile
    labels = {c.GROUP_NAME_LABEL: "kubeflow.org", c.JOB_NAME_LABEL: "tf", c.REPLICA_TYPE_LABEL: "worker",
              c.REPLICA_INDEX_LABEL: "7"}
    cl.api.create({"kind": "Pod", "metadata": {"name": "stray", "namespace": "ns", "labels": labels},
                   "spec": {"containers": [{"name": "tensorflow"}]}})
    cl.kubelet.set_phase("ns", "tf-worker-0", "Running")
    cl.settle()
    stray = cl.api.get("Pod", "ns", "stray")
    assert stray["metadata"]["ownerReferences"][0]["name"] == "tf"
    # index 7 >= replicas: only logged, never deleted (no scale-down, reference pod.go:201-205)
    assert {c.name_of(p) for p in cl.pods("ns")} == {"stray", "tf-worker-0", "tf-worker-1"}


def test_scale_up_creates_missing_indices():
    cl = Cluster()
    cl.submit(_tfjob(workers=1))
    job = cl.job("TFJob", "ns", "tf")
    job["spec"]["tfReplicaSpecs"]["Worker"]["replicas"] = 3
    cl.api.update(job)
    cl.settle()
    assert {c.name_of(p) for p in cl.pods("ns")} == {"tf-worker-0", "tf-worker-1", "tf-worker-2"}


def test_events_emitted():
    cl = Cluster()
    cl.submit(_tfjob(workers=1))
    cl.kubelet.set_phase("ns", "tf-worker-0", "Failed", exit_code=3)
    cl.settle()
    reasons = cl.recs["TFJob"].recorder.reasons()
    for r in ("SuccessfulCreatePod", "SuccessfulCreateService", "ExitedWithCode", "JobFailed"):
        assert r in reasons, reasons
    stored = [e["reason"] for e in cl.api.events_for(cl.job("TFJob", "ns", "tf"))]
    assert "SuccessfulCreatePod" in stored


def test_template_restart_policy_warning():
    cl = Cluster()
    t = tmpl("tensorflow")
    t["spec"]["restartPolicy"] = "Always"
    cl.submit(k.new_job("TFJob", "tf", "ns", {"Worker": {"replicas": 1, "restartPolicy": "Never", "template": t}}))
    assert "SettedPodTemplateRestartPolicy" in cl.recs["TFJob"].recorder.reasons()
    assert cl.pods("ns")[0]["spec"]["restartPolicy"] == "Never"


def test_parallel_pod_creation_slow_start():
    """`--pod-create-parallelism`: same pods, same expectations, created in batches 1, 2, 4, 4...;
    a rejected create stops the ramp and surfaces as the reconcile error."""
    cl = Cluster()
    r = cl.recs["TFJob"]
    r.ctrl.config.pod_create_parallelism = 4
    try:
        batches, lock = [], __import__("threading").Lock()
        real = r.ctrl.create_new_pod

        def spy(job, rt, index, *a, **kw):
            with lock:
                batches.append(int(index))
            return real(job, rt, index, *a, **kw)
        r.ctrl.create_new_pod = spy
        cl.submit(_tfjob(workers=11))
        assert sorted(c.name_of(p) for p in cl.pods("ns")) == sorted(f"tf-worker-{i}" for i in range(11))
        assert sorted(batches) == list(range(11))
        assert batches[0] == 0 and set(batches[1:3]) == {1, 2} and set(batches[3:7]) == {3, 4, 5, 6}   # 1, 2, 4, 4
        assert r.ctrl.expectations.satisfied_expectations(gen_expectation_pods_key("ns/tf", "worker"))
        cl.kubelet.set_all("ns", "Running")
        cl.settle()
        assert cl.job("TFJob", "ns", "tf")["status"]["replicaStatuses"]["Worker"]["active"] == 11
        # failure in the second batch: nothing beyond that batch is attempted in this pass
        del batches[:]
        cl.api.create(_tfjob(name="tf2", workers=8))
        cl.api.fail_next[("create", "Pod")] = RuntimeError("quota exceeded")
        cl.settle()
        assert r.controller.error_count == 1
        cl.mgr.run_until_idle(advance_delayed=True)
        assert len([p for p in cl.pods("ns") if c.name_of(p).startswith("tf2-")]) == 8
    finally:
        r.ctrl.config.pod_create_parallelism = 1
