"""Per-kind controller behaviour: rendezvous env / TF_CONFIG goldens, status rules, restart and
failure paths, plugins.  Golden table from reference controllers/xgboost/pod_test.go:84-137; the
rest covers what the reference left untested."""
import json

import pytest

from cp_helpers import Cluster, load_example, tmpl
from kubedl_b200.api import common as c
from kubedl_b200.api import kinds as k
from kubedl_b200.controllers import tensorflow as tfc
from kubedl_b200.controllers import xdl as xdlc
from kubedl_b200.controllers import xgboost as xgbc


# ---------------------------------------------------------------- PyTorch
def _pt(workers=2, name="pt", **spec):
    return k.new_job("PyTorchJob", name, "ns", {
        "Master": {"replicas": 1, "restartPolicy": "ExitCode", "template": tmpl("pytorch")},
        "Worker": {"replicas": workers, "restartPolicy": "ExitCode", "template": tmpl("pytorch")}}, **spec)


def test_pytorch_env_labels_services():
    cl = Cluster()
    cl.submit(_pt(workers=7))
    pods = {c.name_of(p): p for p in cl.pods("ns")}
    assert len(pods) == 8
    m = cl.env_of(pods["pt-master-0"])
    assert m == {"MASTER_PORT": "23456", "MASTER_ADDR": "localhost", "WORLD_SIZE": "8", "RANK": "0",
                 "PYTHONUNBUFFERED": "0"}
    for i in range(7):
        e = cl.env_of(pods[f"pt-worker-{i}"])
        assert e["MASTER_ADDR"] == "pt-master-0" and e["RANK"] == str(i + 1) and e["WORLD_SIZE"] == "8"
    lab = pods["pt-master-0"]["metadata"]["labels"]
    assert lab == {"group-name": "kubeflow.org", "job-name": "pt", "replica-type": "master",
                   "replica-index": "0", "job-role": "master"}
    assert "job-role" not in pods["pt-worker-0"]["metadata"]["labels"]
    svcs = cl.api.list("Service", "ns")
    assert [c.name_of(s) for s in svcs] == ["pt-master-0"]          # only the Master gets a Service
    assert svcs[0]["spec"]["clusterIP"] == "None"
    assert svcs[0]["spec"]["ports"] == [{"name": "pytorchjob-port", "port": 23456}]
    ref = pods["pt-worker-3"]["metadata"]["ownerReferences"][0]
    assert ref["kind"] == "PyTorchJob" and ref["apiVersion"] == "kubeflow.org/v1" and ref["controller"] \
        and ref["blockOwnerDeletion"]
    assert pods["pt-worker-3"]["spec"]["restartPolicy"] == "Never"   # ExitCode -> Never


def test_pytorch_env_on_every_container_and_custom_port():
    cl = Cluster()
    job = _pt(workers=1)
    job["spec"]["pytorchReplicaSpecs"]["Master"]["template"]["spec"]["containers"] = [
        {"name": "sidecar", "image": "s"},
        {"name": "pytorch", "image": "p", "ports": [{"name": "pytorchjob-port", "containerPort": 31000}]}]
    cl.submit(job)
    master = cl.api.get("Pod", "ns", "pt-master-0")
    for i in range(2):
        assert cl.env_of(master, i)["MASTER_PORT"] == "31000"
    assert cl.env_of(cl.api.get("Pod", "ns", "pt-worker-0"))["MASTER_PORT"] == "31000"


def test_pytorch_lifecycle_success_and_metrics():
    cl = Cluster()
    cl.submit(_pt(workers=2))
    assert cl.conditions("PyTorchJob", "ns", "pt") == [("Created", "True")]
    assert cl.metric("kubedl_jobs_created", kind="pytorchjob") == 1
    assert cl.metric("kubedl_jobs_pending", kind="pytorchjob") == 1
    cl.kubelet.set_all("ns", "Running")
    cl.settle()
    assert cl.conditions("PyTorchJob", "ns", "pt") == [("Created", "True"), ("Running", "True")]
    st = cl.job("PyTorchJob", "ns", "pt")["status"]
    assert st["replicaStatuses"] == {"Master": {"active": 1, "succeeded": 0, "failed": 0},
                                     "Worker": {"active": 2, "succeeded": 0, "failed": 0}}
    assert st["startTime"]
    assert cl.metric("kubedl_jobs_running", kind="pytorchjob") == 1
    assert cl.metric("kubedl_jobs_pending", kind="pytorchjob") == 0
    m = cl.recs["PyTorchJob"].metrics
    assert len(m.launch_delays["first_pod"]) == 1 and len(m.launch_delays["all_pods"]) == 1
    assert cl.metric("kubedl_jobs_all_pods_launch_delay_seconds_count", kind="PyTorchJob", name="pt",
                     namespace="ns", uid=cl.job("PyTorchJob", "ns", "pt")["metadata"]["uid"]) == 1
    # master finishes -> job Succeeded, remaining active folded into succeeded
    cl.kubelet.set_phase("ns", "pt-master-0", "Succeeded")
    cl.settle()
    conds = cl.conditions("PyTorchJob", "ns", "pt")
    assert conds == [("Created", "True"), ("Running", "False"), ("Succeeded", "True")]
    st = cl.job("PyTorchJob", "ns", "pt")["status"]
    assert st["completionTime"] and st["replicaStatuses"]["Worker"] == {"active": 0, "succeeded": 2, "failed": 0}
    assert cl.metric("kubedl_jobs_successful", kind="pytorchjob") == 1
    assert len(cl.pods("ns")) == 3                                    # CleanPodPolicy None (default)
    cl.api.delete("PyTorchJob", "ns", "pt")
    cl.settle()
    assert cl.metric("kubedl_jobs_deleted", kind="pytorchjob") >= 1 and cl.pods("ns") == []


def test_pytorch_exit_code_restart_then_recover():
    cl = Cluster()
    cl.submit(_pt(workers=2))
    cl.kubelet.set_all("ns", "Running")
    cl.settle()
    uid_before = cl.api.get("Pod", "ns", "pt-worker-1")["metadata"]["uid"]
    cl.kubelet.set_phase("ns", "pt-worker-1", "Failed", exit_code=137)     # SIGKILL: retryable
    cl.settle()
    # Restarting is set by the pass that deletes the pod; the pass triggered by the recreated pod
    # sees the master running again and flips back to Running — the transition is in the events
    assert "JobRestarting" in cl.recs["PyTorchJob"].recorder.reasons()
    assert cl.conditions("PyTorchJob", "ns", "pt")[-1] == ("Running", "True")
    recreated = cl.api.get("Pod", "ns", "pt-worker-1")                     # same name/index/rank
    assert recreated["metadata"]["uid"] != uid_before
    assert cl.env_of(recreated)["RANK"] == "2"
    assert cl.metric("kubedl_jobs_restarted", kind="pytorchjob") == 1
    cl.kubelet.set_phase("ns", "pt-worker-1", "Running")
    cl.settle()
    assert cl.conditions("PyTorchJob", "ns", "pt")[-1] == ("Running", "True")
    # all-pods launch delay is NOT re-observed after a restart (case 3, reference job.go:246-259)
    assert len(cl.recs["PyTorchJob"].metrics.launch_delays["all_pods"]) == 1


def test_pytorch_permanent_failure():
    cl = Cluster()
    cl.submit(_pt(workers=1))
    cl.kubelet.set_all("ns", "Running")
    cl.settle()
    cl.kubelet.set_phase("ns", "pt-worker-0", "Failed", exit_code=1)
    cl.settle()
    assert cl.conditions("PyTorchJob", "ns", "pt")[-1] == ("Failed", "True")
    assert cl.metric("kubedl_jobs_failed", kind="pytorchjob") == 1
    assert cl.job("PyTorchJob", "ns", "pt")["status"]["completionTime"]


def test_pytorch_requires_master():
    cl = Cluster()
    cl.api.create(k.new_job("PyTorchJob", "nomaster", "ns",
                            {"Worker": {"replicas": 1, "template": tmpl("pytorch")}}))
    cl.settle()
    assert cl.recs["PyTorchJob"].controller.error_count >= 1     # "failed to find the port" / no master
    assert cl.pods("ns") == []


def test_pytorch_example_yaml_runs_unchanged():
    cl = Cluster()
    (job,) = load_example("example/pytorch/pytorch_job_mnist_mpi.yaml")
    cl.submit(job)
    names = sorted(c.name_of(p) for p in cl.pods("kubedl"))
    assert names == ["pytorch-dist-sendrecv-example-master-0", "pytorch-dist-sendrecv-example-worker-0",
                     "pytorch-dist-sendrecv-example-worker-1"]


# ---------------------------------------------------------------- TensorFlow
def _tf(specs, name="tf", **spec):
    full = {t: {"replicas": n, "restartPolicy": "ExitCode", "template": tmpl("tensorflow")} for t, n in specs.items()}
    return k.new_job("TFJob", name, "ns", full, **spec)


def test_tf_config_golden(monkeypatch):
    monkeypatch.delenv("CUSTOM_CLUSTER_DOMAIN", raising=False)
    cl = Cluster()
    cl.submit(_tf({"PS": 1, "Worker": 2, "Evaluator": 1}))
    w1 = cl.api.get("Pod", "ns", "tf-worker-1")
    cfg = cl.env_of(w1)["TF_CONFIG"]
    assert cfg == ('{"cluster":{"ps":["tf-ps-0.ns.svc:2222"],"worker":["tf-worker-0.ns.svc:2222",'
                   '"tf-worker-1.ns.svc:2222"]},"task":{"type":"worker","index":1},"environment":"cloud"}')
    # Evaluator: accepted, excluded from the cluster map, and never created (reference quirk)
    assert {c.name_of(p) for p in cl.pods("ns")} == {"tf-ps-0", "tf-worker-0", "tf-worker-1"}
    # services for every replica, named like the pods
    assert {c.name_of(s) for s in cl.api.list("Service", "ns")} == {"tf-ps-0", "tf-worker-0", "tf-worker-1"}


def test_tf_config_custom_domain_and_only_tensorflow_container(monkeypatch):
    monkeypatch.setenv("CUSTOM_CLUSTER_DOMAIN", "cluster.local")
    cl = Cluster()
    job = _tf({"Chief": 1, "Worker": 1})
    job["spec"]["tfReplicaSpecs"]["Worker"]["template"]["spec"]["containers"].insert(0, {"name": "sidecar"})
    cl.submit(job)
    w = cl.api.get("Pod", "ns", "tf-worker-0")
    assert "TF_CONFIG" not in cl.env_of(w, 0)
    cfg = json.loads(cl.env_of(w, 1)["TF_CONFIG"])
    assert cfg["cluster"]["chief"] == ["tf-chief-0.ns.svc.cluster.local:2222"]
    assert cl.api.get("Pod", "ns", "tf-chief-0")["metadata"]["labels"]["job-role"] == "master"


def test_tf_single_replica_gets_no_tf_config():
    cl = Cluster()
    (job,) = load_example("example/tf/tf_job_mnist.yaml")
    cl.submit(job)
    (pod,) = cl.pods("kubedl")
    assert "TF_CONFIG" not in cl.env_of(pod)
    assert pod["spec"]["restartPolicy"] == "Never" and pod["spec"]["volumes"][0]["hostPath"]["path"] == "/tmp/data"
    assert tfc.is_distributed(job) is False


def test_tf_status_rules_without_chief():
    cl = Cluster()
    cl.submit(_tf({"PS": 1, "Worker": 2}))
    cl.kubelet.set_all("ns", "Running")
    cl.settle()
    assert cl.conditions("TFJob", "ns", "tf")[-1] == ("Running", "True")
    cl.kubelet.set_phase("ns", "tf-worker-0", "Succeeded", exit_code=0)    # worker-0 done -> job done
    cl.settle()
    assert cl.conditions("TFJob", "ns", "tf")[-1] == ("Succeeded", "True")
    # CleanPodPolicy Running (TF default): the still-running PS and worker-1 are removed
    assert {c.name_of(p) for p in cl.pods("ns")} == {"tf-worker-0"}


def test_tf_status_rules_with_chief():
    cl = Cluster()
    cl.submit(_tf({"Chief": 1, "Worker": 2}))
    cl.kubelet.set_all("ns", "Running")
    cl.settle()
    cl.kubelet.set_phase("ns", "tf-worker-0", "Succeeded")
    cl.kubelet.set_phase("ns", "tf-worker-1", "Succeeded")
    cl.settle()
    assert cl.conditions("TFJob", "ns", "tf")[-1] == ("Running", "True")   # chief decides
    cl.kubelet.set_phase("ns", "tf-chief-0", "Succeeded")
    cl.settle()
    assert cl.conditions("TFJob", "ns", "tf")[-1] == ("Succeeded", "True")


# ---------------------------------------------------------------- XGBoost
@pytest.mark.parametrize("workers,rtype,index,expect", [
    (1, "master", 0, {"WORLD_SIZE": "2", "RANK": "0"}), (1, "worker", 0, {"WORLD_SIZE": "2", "RANK": "0"}),
    (2, "worker", 1, {"WORLD_SIZE": "3", "RANK": "1"}), (7, "worker", 6, {"WORLD_SIZE": "8", "RANK": "6"})])
def test_xgboost_set_pod_env_table(workers, rtype, index, expect):
    job = k.new_job("XGBoostJob", "xgb", "ns", {
        "Master": {"replicas": 1, "template": tmpl("xgboostjob", ports=[{"name": "xgboostjob-port", "containerPort": 9999}])},
        "Worker": {"replicas": workers, "template": tmpl("xgboostjob")}})
    t = tmpl("xgboostjob")
    xgbc.set_pod_env(job, t, str(index))
    env = {e["name"]: e["value"] for e in t["spec"]["containers"][0]["env"]}
    assert env["MASTER_PORT"] == "9999" and env["MASTER_ADDR"] == "xgb-master-0" and env["PYTHONUNBUFFERED"] == "0"
    for key, v in expect.items():
        assert env[key] == v


def test_xgboost_example_and_lifecycle():
    cl = Cluster()
    (job,) = load_example("example/xgboost/xgboostjob_v1alpha1_iris_train.yaml")
    cl.submit(job)
    pods = {c.name_of(p): p for p in cl.pods("default")}
    assert set(pods) == {"xgboost-dist-iris-test-train-master-0", "xgboost-dist-iris-test-train-worker-0",
                         "xgboost-dist-iris-test-train-worker-1"}
    e = cl.env_of(pods["xgboost-dist-iris-test-train-worker-1"])
    assert e["MASTER_PORT"] == "9991" and e["RANK"] == "1" and e["WORLD_SIZE"] == "3"
    assert e["MASTER_ADDR"] == "xgboost-dist-iris-test-train-master-0"
    assert pods["xgboost-dist-iris-test-train-master-0"]["spec"]["restartPolicy"] == "Never"
    assert len(cl.api.list("Service", "default")) == 3
    j = cl.job("XGBoostJob", "default", "xgboost-dist-iris-test-train")
    assert j["metadata"]["labels"] if False else True
    assert pods["xgboost-dist-iris-test-train-worker-0"]["metadata"]["labels"]["group-name"] == "kubeflow.org"
    # Running is asserted at the end of every non-terminal pass, even before pods run
    assert cl.conditions("XGBoostJob", "default", "xgboost-dist-iris-test-train")[-1] == ("Running", "True")
    cl.kubelet.set_all("default", "Running")
    cl.settle()
    assert cl.job("XGBoostJob", "default", "xgboost-dist-iris-test-train")["status"]["startTime"]
    cl.kubelet.set_phase("default", "xgboost-dist-iris-test-train-master-0", "Succeeded")
    cl.settle()
    assert cl.conditions("XGBoostJob", "default", "xgboost-dist-iris-test-train")[-1] == ("Succeeded", "True")
    # default TTL of 100 s: requeued, not deleted yet
    assert cl.recs["XGBoostJob"].controller.queue.pending_delayed() >= 1


def test_xgboost_empty_restart_policy_passthrough():
    cl = Cluster()
    cl.submit(k.new_job("XGBoostJob", "x", "ns", {"Master": {"template": tmpl("xgboostjob")}}))
    assert cl.api.get("Pod", "ns", "x-master-0")["spec"]["restartPolicy"] == ""


# ---------------------------------------------------------------- XDL
def test_xdl_example_env_and_zk_addr():
    cl = Cluster()
    docs = load_example("example/xdl/xdl_job_mnist.yaml")
    job = [d for d in docs if d["kind"] == "XDLJob"][0]
    created = cl.submit(job)
    uid = created["metadata"]["uid"]
    pods = {c.name_of(p): p for p in cl.pods("kubedl")}
    assert set(pods) == {"xdl-mnist-example-ps-0", "xdl-mnist-example-scheduler-0",
                         "xdl-mnist-example-worker-0", "xdl-mnist-example-worker-1"}
    e = cl.env_of(pods["xdl-mnist-example-worker-1"])
    assert e == {"ZK_ADDR": f"zfs://zk-0:2181/xdl/{uid}", "TASK_NAME": "worker", "TASK_INDEX": "1"}
    assert "job-role" not in pods["xdl-mnist-example-scheduler-0"]["metadata"]["labels"]
    assert pods["xdl-mnist-example-ps-0"]["metadata"]["labels"]["group-name"] == "xdl.kubedl.io"
    assert cl.job("XDLJob", "kubedl", "xdl-mnist-example")["spec"]["ttlSecondsAfterFinished"] == 3600


@pytest.mark.parametrize("spec,workers,expect", [({"minFinishWorkRate": 90}, 10, 9), ({"minFinishWorkRate": 50}, 3, 2),
                                                 ({"minFinishWorkNum": 4}, 10, 4), ({}, 7, 7),
                                                 ({"minFinishWorkRate": 10, "minFinishWorkNum": 9}, 10, 1)])
def test_xdl_calculate_min_finish(spec, workers, expect):
    assert xdlc.calculate_min_finish({"spec": spec}, workers) == expect


def test_xdl_partial_success_and_failure():
    specs = {"PS": {"replicas": 1, "restartPolicy": "Never", "template": tmpl("xdl")},
             "Worker": {"replicas": 4, "restartPolicy": "ExitCode", "template": tmpl("xdl")}}
    cl = Cluster()
    cl.submit(k.new_job("XDLJob", "x", "ns", specs, minFinishWorkRate=50))
    cl.kubelet.set_all("ns", "Running")
    cl.settle()
    assert cl.conditions("XDLJob", "ns", "x")[-1] == ("Running", "True")
    cl.kubelet.set_phase("ns", "x-worker-0", "Succeeded")
    cl.settle()
    assert cl.conditions("XDLJob", "ns", "x")[-1] == ("Running", "True")
    cl.kubelet.set_phase("ns", "x-worker-1", "Succeeded")                 # 2 of 4 = 50 %
    cl.settle()
    assert cl.conditions("XDLJob", "ns", "x")[-1] == ("Succeeded", "True")
    cl2 = Cluster()
    cl2.submit(k.new_job("XDLJob", "x", "ns", specs))
    cl2.kubelet.set_all("ns", "Running")
    cl2.settle()
    cl2.kubelet.set_phase("ns", "x-ps-0", "Failed", exit_code=1)
    cl2.settle()
    assert cl2.conditions("XDLJob", "ns", "x")[-1] == ("Failed", "True")
    assert "XdlJobFailed" in cl2.recs["XDLJob"].recorder.reasons()


def test_xdl_config_generator():
    job = k.new_job("XDLJob", "x", "ns", {"PS": {"replicas": 1, "template": tmpl("xdl")},
                                          "Worker": {"replicas": 2, "template": tmpl("xdl")}})
    k.SCHEME.default(job)
    cfg = json.loads(xdlc.gen_xdl_config(job, "worker", "1"))
    assert cfg == {"cluster": {"ps": ["x-ps-0:2222"], "worker": ["x-worker-0:2222", "x-worker-1:2222"]},
                   "task": {"type": "worker", "index": 1}}


# ---------------------------------------------------------------- plugins
def test_gang_scheduling_kube_batch():
    cl = Cluster(gang="kube-batch")
    cl.submit(_pt(workers=3, cleanPodPolicy="All"))
    pg = cl.api.get("PodGroup", "ns", "pt")
    assert pg["spec"]["minMember"] == 4 and pg["metadata"]["ownerReferences"][0]["kind"] == "PyTorchJob"
    assert all(p["spec"]["schedulerName"] == "kube-batch" for p in cl.pods("ns"))
    cl.kubelet.set_all("ns", "Running")
    cl.settle()
    cl.kubelet.set_phase("ns", "pt-master-0", "Succeeded")
    cl.settle()
    assert cl.api.list("PodGroup", "ns") == []                            # deleted on terminal state
    reasons = cl.recs["PyTorchJob"].recorder.reasons()
    assert "JobTerminated" in reasons and "SuccessfulDeletePodGroup" in reasons


def test_gpu_gang_honours_min_available():
    """schedulingPolicy.minAvailable is declared but unused by the reference; the gpu-gang scheduler uses
    it as the gang size (capped by the replica count)."""
    cl = Cluster(gang="gpu-gang")
    job = _pt(workers=3)
    job["spec"]["schedulingPolicy"] = {"minAvailable": 2}
    cl.submit(job)
    pg = cl.api.get("PodGroup", "ns", "pt")
    assert pg["spec"]["minMember"] == 2
    assert all(p["spec"]["schedulerName"] == "gpu-gang" for p in cl.pods("ns"))
    assert all(p["metadata"]["annotations"]["kubedl.io/gang-name"] == "pt" for p in cl.pods("ns"))


def test_gang_scheduler_registry_and_unknown_name():
    from kubedl_b200.plugins import gang_schedule
    cl = Cluster()
    assert set(gang_schedule.REGISTRY.names()) >= {"kube-batch", "gpu-gang"}
    with pytest.raises(ValueError):
        Cluster(gang="no-such-scheduler")


def test_code_sync_injection():
    cl = Cluster()
    job = _pt(workers=1)
    job["metadata"]["annotations"] = {c.ANNOTATION_GIT_SYNC_CONFIG: json.dumps(
        {"source": "https://github.com/sample/sample.git", "branch": "dev", "user": "u", "password": "p"})}
    job["spec"]["pytorchReplicaSpecs"]["Worker"]["template"]["spec"]["containers"][0]["workingDir"] = "/work"
    job["spec"]["pytorchReplicaSpecs"]["Worker"]["template"]["spec"]["containers"][0]["resources"] = {
        "limits": {"cpu": "2"}}
    cl.submit(job)
    w = cl.api.get("Pod", "ns", "pt-worker-0")
    (ic,) = w["spec"]["initContainers"]
    assert ic["name"] == "git-sync-code" and ic["image"] == "kubedl/git-sync:v1"
    env = {e["name"]: e["value"] for e in ic["env"]}
    assert env == {"GIT_SYNC_REPO": "https://github.com/sample/sample.git", "GIT_SYNC_ONE_TIME": "true",
                   "GIT_SYNC_MAX_SYNC_FAILURES": "3", "GIT_SYNC_BRANCH": "dev", "GIT_SYNC_ROOT": "/code",
                   "GIT_SYNC_DEST": "sample", "GIT_SYNC_USERNAME": "u", "GIT_SYNC_PASSWORD": "p"}
    assert ic["resources"] == {"limits": {"cpu": "2"}}                     # copies Containers[0] resources
    assert w["spec"]["volumes"] == [{"name": "git-sync", "emptyDir": {}}]
    assert w["spec"]["containers"][0]["volumeMounts"] == [
        {"name": "git-sync", "readOnly": False, "mountPath": "/work/sample", "subPath": "sample"}]
    assert len(cl.api.get("Pod", "ns", "pt-master-0")["spec"]["initContainers"]) == 1


def test_workload_gate_disables_controller():
    cl = Cluster(workloads="PyTorchJob,TFJob")
    assert set(cl.recs) == {"PyTorchJob", "TFJob"}
    cl.submit(k.new_job("XDLJob", "x", "ns", {"Worker": {"template": tmpl("xdl")}}))
    assert cl.pods("ns") == []


def test_status_schema_matches_crd_requirements():
    """CRD validation requires status.conditions and status.replicaStatuses
    (reference config/crd/bases/kubeflow.org_pytorchjobs.yaml:149-151)."""
    cl = Cluster()
    cl.submit(_pt(workers=1))
    st = cl.job("PyTorchJob", "ns", "pt")["status"]
    assert isinstance(st["conditions"], list) and isinstance(st["replicaStatuses"], dict)
    cond = st["conditions"][0]
    assert set(cond) == {"type", "status", "reason", "message", "lastUpdateTime", "lastTransitionTime"}


# ---------------------------------------------------------------- status writes are compare-and-swap
def test_status_write_from_a_stale_job_is_refused_and_requeued():
    """Status().Update carries the resourceVersion of the object the status was computed from (reference
    controllers/pytorch/job.go:95-105): a write derived from a stale read gets Conflict instead of
    overwriting what is newer, and the reconcile is simply run again."""
    from kubedl_b200.apiserver.store import Conflict
    cl = Cluster()
    cl.submit(_pt(workers=1))
    rec = cl.recs["PyTorchJob"]
    stale = cl.job("PyTorchJob", "ns", "pt")
    # somebody else (a second manager) moves the job on
    newer = cl.job("PyTorchJob", "ns", "pt")
    newer["status"]["startTime"] = "2026-01-01T00:00:00Z"
    cl.api.update_status(newer)
    lost = c.deepcopy(stale["status"])
    with pytest.raises(Conflict):
        rec.update_job_status_in_api_server(stale, lost)
    assert cl.job("PyTorchJob", "ns", "pt")["status"]["startTime"] == "2026-01-01T00:00:00Z"
    # through the controller: the conflict is a quiet requeue, and the next pass converges
    ctl = rec.controller
    errors0 = ctl.error_count
    real, calls = cl.mgr.client.update_status, []

    def flaky(obj):
        calls.append(obj["metadata"].get("resourceVersion"))
        if len(calls) == 1:
            raise Conflict("the object has been modified")
        return real(obj)
    cl.mgr.client.update_status = flaky
    try:
        cl.kubelet.set_all("ns", "Running")
        cl.mgr.run_until_idle(advance_delayed=True)
    finally:
        cl.mgr.client.update_status = real
    assert len(calls) >= 2 and all(calls), calls          # every status write named a resourceVersion
    assert ("Running", "True") in cl.conditions("PyTorchJob", "ns", "pt")
    assert ctl.error_count == errors0 and ctl.conflict_count >= 1


def test_created_condition_is_added_not_overwritten():
    """The Created stamp of the watch thread goes on top of whatever status the API server holds."""
    cl = Cluster()
    job = _pt(workers=1)
    job["status"] = {"conditions": [], "replicaStatuses": {}, "startTime": "2026-01-01T00:00:00Z"}
    cl.api.create(c.deepcopy(job))
    cl.settle()
    st_ = cl.job("PyTorchJob", "ns", "pt")["status"]
    assert st_["conditions"][0]["type"] == "Created" and st_.get("startTime") == "2026-01-01T00:00:00Z"


def test_no_watcher_writes_to_the_shared_event_snapshot(monkeypatch):
    """Watchers registered with shared=True (informer cache, controllers, node agent) are handed ONE snapshot per event.
    With the snapshots frozen, a job of every kind goes through create -> pods Running -> Succeeded without any handler
    trying to write to what it was handed (on_owner_create used to default the event's job in place)."""
    from kubedl_b200.apiserver import store
    monkeypatch.setattr(store, "_FREEZE_SNAPSHOTS", True)
    from kubedl_b200.bench.launch_delay import _job
    cl = Cluster()
    for kind in ("PyTorchJob", "TFJob", "XGBoostJob", "XDLJob"):
        name = "frozen-" + kind.lower()
        cl.submit(_job(kind, name, 3, 0.0))
        assert ("Created", "True") in cl.conditions(kind, "bench", name)
        cl.kubelet.set_all("bench", c.POD_RUNNING, {c.JOB_NAME_LABEL: name})
        cl.settle()
        cl.kubelet.set_all("bench", c.POD_SUCCEEDED, {c.JOB_NAME_LABEL: name})
        cl.settle()
        assert cl.conditions(kind, "bench", name)[-1] == ("Succeeded", "True"), cl.conditions(kind, "bench", name)
    assert sum(r.controller.error_count for r in cl.recs.values()) == 0


def test_informer_cache_lists_a_jobs_objects_through_its_index():
    """Listing by {group-name, job-name} goes through the cache's job-name index; the index follows label changes
    (adoption / release) and deletions, and the result is what a scan of the whole cache gives."""
    cl = Cluster()
    api, client = cl.api, cl.mgr.client
    client.ensure_informer("Pod")

    def mk(name, job):
        labels = {c.JOB_NAME_LABEL: job, c.GROUP_NAME_LABEL: "kubeflow.org"} if job else {"other": "x"}
        return {"apiVersion": "v1", "kind": "Pod", "metadata": {"name": name, "namespace": "ns", "labels": labels},
                "spec": {"containers": [{"name": "c", "image": "i"}]}}
    for i in range(3):
        api.create(mk(f"a-{i}", "job-a"))
    api.create(mk("b-0", "job-b"))
    api.create(mk("loose", ""))

    def names(job):
        return [c.name_of(p) for p in client.list("Pod", "ns", {c.JOB_NAME_LABEL: job, c.GROUP_NAME_LABEL: "kubeflow.org"})]

    def scan(job):
        return sorted(c.name_of(p) for p in client.list("Pod", "ns") if c.labels_of(p).get(c.JOB_NAME_LABEL) == job)
    assert names("job-a") == ["a-0", "a-1", "a-2"] == scan("job-a") and names("job-b") == ["b-0"] and names("nope") == []
    api.patch_labels("Pod", "ns", "a-1", {c.JOB_NAME_LABEL: "job-b"})          # moves between jobs
    api.patch_labels("Pod", "ns", "loose", {c.JOB_NAME_LABEL: "job-a", c.GROUP_NAME_LABEL: "kubeflow.org"})   # adopted
    api.patch_labels("Pod", "ns", "a-2", {c.JOB_NAME_LABEL: None})             # released
    api.delete("Pod", "ns", "a-0")
    assert names("job-a") == ["loose"] == scan("job-a")
    assert names("job-b") == ["a-1", "b-0"] == scan("job-b")
    assert client.list("Pod", "other-ns", {c.JOB_NAME_LABEL: "job-b"}) == []
    assert not any(not keys for keys in client._by_job["Pod"].values())          # empty buckets are dropped


def test_nothing_is_kept_about_a_deleted_job():
    """A job that ran and was deleted leaves no expectation record, no back-off state, no cache entry and no index
    bucket behind (the reference's expectations live in a TTL store; a plain dict kept 3-4 records per job for ever)."""
    from kubedl_b200.bench.launch_delay import _job
    cl = Cluster()
    for kind in ("PyTorchJob", "TFJob", "XGBoostJob", "XDLJob"):
        for i in range(3):
            cl.api.create(_job(kind, f"gone-{kind.lower()}-{i}", 3, 0.0))
    cl.settle()
    assert sum(len(r.ctrl.expectations) for r in cl.recs.values()) > 0
    cl.kubelet.set_all("bench", c.POD_RUNNING); cl.settle()
    cl.kubelet.set_all("bench", c.POD_SUCCEEDED); cl.settle()
    for kind in ("PyTorchJob", "TFJob", "XGBoostJob", "XDLJob"):
        for i in range(3):
            try:                                        # XGBoostJobs may already be gone (TTL after finished)
                cl.api.delete(kind, "bench", f"gone-{kind.lower()}-{i}")
            except Exception:
                pass
    cl.settle()
    assert [len(r.ctrl.expectations) for r in cl.recs.values()] == [0, 0, 0, 0]
    assert all(len(v) == 0 for kind, v in cl.mgr.client._cache.items() if kind in ("Pod", "Service") or kind.endswith("Job"))
    assert all(len(v) == 0 for v in cl.mgr.client._by_job.values())


def test_expectations_sweep_drops_expired_records():
    from kubedl_b200.runtime.expectations import ControllerExpectations
    now = [0.0]
    e = ControllerExpectations(clock=lambda: now[0], ttl=300.0)
    e.expect_creations("ns/a/worker/pods", 2)
    e.expect_creations("ns/b/worker/pods", 1)
    now[0] = 200.0
    e.expect_creations("ns/c/worker/pods", 1)
    now[0] = 400.0
    assert e.sweep() == 2 and len(e) == 1 and e.get("ns/c/worker/pods") is not None
    assert e.delete_for_job("ns/c") == 1 and len(e) == 0
