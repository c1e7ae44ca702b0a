"""API types: defaulting tables and storage round trips.
Ports the ideas of reference api/tensorflow/v1/defaults_test.go:78-276,
api/xdl/v1alpha1/defaults_test.go:91-511, api/*/types_test.go, api/tensorflow/v1/util_test.go."""
import copy

import pytest

from kubedl_b200.api import common as c
from kubedl_b200.api import kinds as k
from kubedl_b200.api.kinds import SCHEME
from kubedl_b200.apiserver.store import AlreadyExists, APIServer, Conflict, NotFound


def _tmpl(container="tensorflow", image="img", ports=None):
    ctr = {"name": container, "image": image}
    if ports is not None:
        ctr["ports"] = ports
    return {"spec": {"containers": [ctr]}}


def _tf(specs, **spec):
    return {"apiVersion": "kubeflow.org/v1", "kind": "TFJob", "metadata": {"name": "j", "namespace": "ns"},
            "spec": {"tfReplicaSpecs": specs, **spec}}


def test_tf_set_replicas_and_case_normalisation():
    job = _tf({"worker": {"template": _tmpl(ports=[{"name": "tfjob-port", "containerPort": 2222}])}})
    SCHEME.default(job)
    specs = job["spec"]["tfReplicaSpecs"]
    assert list(specs) == ["Worker"]
    assert specs["Worker"]["replicas"] == 1
    assert specs["Worker"]["restartPolicy"] == "ExitCode"   # "set replicas with default restartpolicy"
    assert job["spec"]["cleanPodPolicy"] == "Running"


def test_tf_set_replicas_with_default_port():
    job = _tf({"Worker": {"replicas": 3, "restartPolicy": "OnFailure", "template": _tmpl()}})
    SCHEME.default(job)
    w = job["spec"]["tfReplicaSpecs"]["Worker"]
    assert w["replicas"] == 3 and w["restartPolicy"] == "OnFailure"
    assert w["template"]["spec"]["containers"][0]["ports"] == [{"name": "tfjob-port", "containerPort": 2222}]


def test_tf_set_replicas_adding_default_port():
    job = _tf({"PS": {"template": _tmpl(ports=[{"name": "custom", "containerPort": 9999}])}})
    SCHEME.default(job)
    ports = job["spec"]["tfReplicaSpecs"]["PS"]["template"]["spec"]["containers"][0]["ports"]
    assert ports == [{"name": "custom", "containerPort": 9999}, {"name": "tfjob-port", "containerPort": 2222}]


def test_tf_custom_port_kept_and_custom_cleanpod_policy():
    job = _tf({"Worker": {"template": _tmpl(ports=[{"name": "tfjob-port", "containerPort": 7777}])}},
              cleanPodPolicy="All")
    SCHEME.default(job)
    assert job["spec"]["cleanPodPolicy"] == "All"
    ports = job["spec"]["tfReplicaSpecs"]["Worker"]["template"]["spec"]["containers"][0]["ports"]
    assert ports == [{"name": "tfjob-port", "containerPort": 7777}]


def test_default_port_goes_to_named_container_else_first():
    spec = {"containers": [{"name": "sidecar"}, {"name": "tensorflow"}]}
    k.set_default_port(spec, "tensorflow", "tfjob-port", 2222)
    assert "ports" not in spec["containers"][0] and spec["containers"][1]["ports"][0]["containerPort"] == 2222
    spec = {"containers": [{"name": "a"}, {"name": "b"}]}
    k.set_default_port(spec, "tensorflow", "tfjob-port", 2222)
    assert spec["containers"][0]["ports"][0]["name"] == "tfjob-port"


def test_tf_helpers():
    assert k.tf_is_chief_or_master("Chief") and k.tf_is_chief_or_master("Master")
    assert not k.tf_is_chief_or_master("Worker")
    assert k.tf_is_worker("Worker") and k.tf_is_evaluator("Evaluator")


def test_pytorch_defaults():
    job = {"kind": "PyTorchJob", "metadata": {"name": "p"},
           "spec": {"pytorchReplicaSpecs": {"master": {"template": _tmpl("pytorch")},
                                            "WORKER": {"replicas": 4, "template": _tmpl("pytorch")}}}}
    SCHEME.default(job)
    s = job["spec"]
    assert s["cleanPodPolicy"] == "None"
    assert set(s["pytorchReplicaSpecs"]) == {"Master", "Worker"}
    m, w = s["pytorchReplicaSpecs"]["Master"], s["pytorchReplicaSpecs"]["Worker"]
    assert m["replicas"] == 1 and m["restartPolicy"] == "ExitCode"
    assert w["replicas"] == 4 and w["restartPolicy"] == "OnFailure"
    assert m["template"]["spec"]["containers"][0]["ports"] == [{"name": "pytorchjob-port", "containerPort": 23456}]
    assert "ports" not in w["template"]["spec"]["containers"][0]   # default port only on Master


def test_xgboost_defaults():
    job = {"kind": "XGBoostJob", "metadata": {"name": "x"},
           "spec": {"xgbReplicaSpecs": {"Master": {"template": _tmpl("xgboostjob")},
                                        "worker": {"template": _tmpl("xgboostjob")}}}}
    SCHEME.default(job)
    s = job["spec"]
    assert s["cleanPodPolicy"] == "None" and s["ttlSecondsAfterFinished"] == 100
    for rt in ("Master", "Worker"):
        rs = s["xgbReplicaSpecs"][rt]
        assert rs["replicas"] == 1 and "restartPolicy" not in rs       # no default restart policy
        assert rs["template"]["spec"]["containers"][0]["ports"] == [{"name": "xgboostjob-port", "containerPort": 9999}]


def _xdl(specs=None, **spec):
    return {"kind": "XDLJob", "metadata": {"name": "x"},
            "spec": {"xdlReplicaSpecs": specs or {"Worker": {"template": _tmpl("xdl")}}, **spec}}


def test_xdl_default_min_finish_attributes():
    job = _xdl()
    SCHEME.default(job)
    s = job["spec"]
    assert s["minFinishWorkRate"] == 90 and "minFinishWorkNum" not in s
    assert s["backoffLimit"] == 20 and s["cleanPodPolicy"] == "Running"
    w = s["xdlReplicaSpecs"]["Worker"]
    assert w["restartPolicy"] == "Never" and w["replicas"] == 1
    assert w["template"]["spec"]["containers"][0]["ports"] == [{"name": "xdljob-port", "containerPort": 2222}]


def test_xdl_set_min_finish_work_num_percentage_backoff():
    job = _xdl(minFinishWorkNum=3)
    SCHEME.default(job)
    assert job["spec"]["minFinishWorkNum"] == 3 and "minFinishWorkRate" not in job["spec"]
    job = _xdl(minFinishWorkRate=50)
    SCHEME.default(job)
    assert job["spec"]["minFinishWorkRate"] == 50
    job = _xdl(backoffLimit=5)
    SCHEME.default(job)
    assert job["spec"]["backoffLimit"] == 5


def test_xdl_case_normalisation_all_types():
    job = _xdl({"ps": {"template": _tmpl("xdl")}, "scheduler": {"template": _tmpl("xdl")},
                "extendrole": {"template": _tmpl("xdl")}, "worker": {"template": _tmpl("xdl")}})
    SCHEME.default(job)
    assert set(job["spec"]["xdlReplicaSpecs"]) == {"PS", "Scheduler", "ExtendRole", "Worker"}


def test_defaulting_is_idempotent():
    job = _tf({"worker": {"template": _tmpl()}, "ps": {"template": _tmpl()}})
    SCHEME.default(job)
    once = copy.deepcopy(job)
    SCHEME.default(job)
    assert job == once


@pytest.mark.parametrize("kind", ["PyTorchJob", "TFJob", "XGBoostJob", "XDLJob"])
def test_storage_round_trip(kind):
    """Create / Get / Update labels / Delete of each kind (reference api/pytorch/v1/types_test.go:30-88)."""
    api = APIServer()
    info = SCHEME.info(kind)
    job = k.new_job(kind, "foo", "default", {info.replica_types[0]: {"template": _tmpl(info.default_container_name)}})
    created = api.create(job)
    assert created["metadata"]["uid"] and created["metadata"]["resourceVersion"]
    assert created["apiVersion"] == info.api_version
    with pytest.raises(AlreadyExists):
        api.create(job)
    fetched = api.get(kind, "default", "foo")
    assert fetched == created
    fetched["metadata"]["labels"] = {"hello": "world"}
    updated = api.update(fetched)
    assert api.get(kind, "default", "foo")["metadata"]["labels"] == {"hello": "world"}
    # stale resourceVersion -> Conflict
    with pytest.raises(Conflict):
        api.update(fetched)
    # status subresource: update() ignores status, update_status() ignores spec
    u2 = copy.deepcopy(updated)
    u2["status"] = {"conditions": [{"type": "Created"}]}
    u2 = api.update(u2)
    assert "status" not in api.get(kind, "default", "foo")
    u2["status"] = {"conditions": [{"type": "Created"}], "replicaStatuses": {}}
    u2["spec"]["cleanPodPolicy"] = "All"
    api.update_status(u2)
    got = api.get(kind, "default", "foo")
    assert got["status"]["conditions"][0]["type"] == "Created" and got["spec"].get("cleanPodPolicy") != "All"
    api.delete(kind, "default", "foo")
    with pytest.raises(NotFound):
        api.get(kind, "default", "foo")


def test_owner_reference_garbage_collection_and_watch():
    api = APIServer()
    events = []
    api.watch("Pod", lambda ev, obj, old: events.append((ev, c.name_of(obj))))
    job = api.create(k.new_job("TFJob", "j", "ns", {"Worker": {"template": _tmpl()}}))
    ref = {"apiVersion": job["apiVersion"], "kind": "TFJob", "name": "j", "uid": job["metadata"]["uid"],
           "controller": True, "blockOwnerDeletion": True}
    api.create({"kind": "Pod", "metadata": {"name": "p0", "namespace": "ns", "ownerReferences": [ref],
                                             "labels": {"a": "b"}}, "spec": {}})
    api.create({"kind": "Pod", "metadata": {"name": "other", "namespace": "ns"}, "spec": {}})
    assert [c.name_of(p) for p in api.list("Pod", "ns", {"a": "b"})] == ["p0"]
    api.delete("TFJob", "ns", "j")
    assert [c.name_of(p) for p in api.list("Pod", "ns")] == ["other"]
    assert ("ADDED", "p0") in events and ("DELETED", "p0") in events


def test_crd_discovery_gates_kinds():
    api = APIServer(install_all_crds=False)
    assert not api.crd_installed("TFJob")
    with pytest.raises(NotFound):
        api.create(k.new_job("TFJob", "j"))
    api.install_crd("TFJob")
    assert api.server_resources_for_group_version("kubeflow.org/v1") == ["TFJob"]
    api.create(k.new_job("TFJob", "j"))


def test_events_expire_like_kube_apiserver():
    """--event-ttl: Events older than an hour are dropped so a long-running manager does not grow."""
    import time
    from kubedl_b200.apiserver.store import APIServer
    api = APIServer()
    job = api.create(k.new_job("TFJob", "e", "ns", replica_specs={}))
    for i in range(3):
        api.record_event(job, "Normal", "SuccessfulCreatePod", f"pod {i}")
    assert len(api.events_for(job)) == 3
    assert api.prune_events(now=time.time() + 10) == 0
    assert api.prune_events(now=time.time() + 3700) == 3 and api.events_for(job) == []
    assert api.request_log.maxlen == 65536


def test_a_failing_watch_handler_costs_neither_the_writer_nor_the_other_watchers(caplog):
    """A watcher that raises is logged and skipped: the create that caused the event returns normally and the watchers
    registered after it still see the event (the remote client's reflector already behaves this way)."""
    from kubedl_b200.apiserver.store import APIServer
    api = APIServer()
    seen = []

    def bad(ev, obj, old):
        raise RuntimeError("boom")
    api.watch("Pod", bad, replay=False)
    api.watch("Pod", lambda ev, obj, old: seen.append((ev, obj["metadata"]["name"])), replay=False)
    out = api.create({"apiVersion": "v1", "kind": "Pod", "metadata": {"name": "p", "namespace": "ns"},
                      "spec": {"containers": [{"name": "c", "image": "i"}]}})
    assert out["metadata"]["name"] == "p" and seen == [("ADDED", "p")]
    assert "watch handler for Pod failed" in caplog.text and "boom" in caplog.text
