"""The test-fixture package (reference pkg/test_job, pkg/test_util) driving the engine directly, the
generated CRDs, and the custom-workload recipe (reference docs/how-to-add-a-custom-workload.md)."""
import os

import pytest
import yaml

from kubedl_b200 import testing as T
from kubedl_b200.api import common as c
from kubedl_b200.api import crd
from kubedl_b200.api.kinds import SCHEME
from kubedl_b200.apiserver.store import APIServer
from kubedl_b200.engine.job_controller import JobController, JobControllerConfiguration
from kubedl_b200.runtime.manager import EventRecorder, Manager


def _engine():
    mgr = Manager()
    ctl = T.TestJobController()
    jc = JobController(ctl, JobControllerConfiguration(), mgr.client, EventRecorder(mgr.api, "test"))
    return jc, ctl


@pytest.mark.parametrize("policy,deleted_pods", [("Running", {"worker-1"}), ("All", {"worker-0", "worker-1", "worker-2"}),
                                                 ("None", set())])
def test_delete_pods_and_services_with_stub_controller(policy, deleted_pods):
    """reference pkg/job_controller/job_test.go:15-110"""
    jc, ctl = _engine()
    job = T.new_test_job(3)
    T.set_pods_statuses(ctl.pods, job, "worker", pending=1, active=1, succeeded=1, failed=0)
    ctl.services = [T.new_service(job, "worker", i) for i in range(3)]
    jc.delete_pods_and_services({"cleanPodPolicy": policy}, job, list(ctl.pods))
    assert {c.name_of(p) for p in ctl.pods} == {"worker-0", "worker-1", "worker-2"} - deleted_pods
    assert {c.name_of(s) for s in ctl.services} == {"worker-0", "worker-1", "worker-2"} - deleted_pods


def test_reconcile_with_stub_controller_creates_children_and_ttl_deletes():
    jc, ctl = _engine()
    job = T.new_test_job(2)
    ctl.job = job
    specs = job["spec"]["testReplicaSpecs"]
    jc.reconcile_jobs(job, specs, job["status"], {"cleanPodPolicy": "None"})
    assert sorted(c.name_of(p) for p in ctl.pods) == ["test-job-worker-0", "test-job-worker-1"]
    assert ctl.pods[0]["metadata"]["labels"][c.GROUP_NAME_LABEL] == T.LABEL_GROUP_NAME
    assert sorted(c.name_of(s) for s in ctl.services) == ["test-job-worker-0", "test-job-worker-1"]
    # finished + TTL 0 -> cleanupJob deletes it (reference job_test.go:186-216)
    job["status"]["completionTime"] = c.now_rfc3339()
    from kubedl_b200.utils import status as st
    st.update_job_conditions(job["status"], c.JOB_SUCCEEDED, st.JOB_SUCCEEDED_REASON, "done")
    assert T.check_condition(job, c.JOB_SUCCEEDED, st.JOB_SUCCEEDED_REASON)
    jc.reconcile_jobs(job, specs, job["status"], {"cleanPodPolicy": "None", "ttlSecondsAfterFinished": 0})
    assert ctl.job is None


def test_crd_manifests_match_registry(tmp_path):
    paths = crd.write_all(str(tmp_path))
    assert sorted(os.path.basename(p) for p in paths) == [
        "kubeflow.org_pytorchjobs.yaml", "kubeflow.org_tfjobs.yaml", "xdl.kubedl.io_xdljobs.yaml",
        "xgboostjob.kubeflow.org_xgboostjobs.yaml"]
    for p in paths:
        doc = list(yaml.safe_load_all(open(p)))[-1]
        committed = list(yaml.safe_load_all(open(os.path.join("config/crd/bases", os.path.basename(p)))))[-1]
        assert doc == committed, f"{p}: run python -m kubedl_b200.api.crd"
        spec = doc["spec"]
        assert spec["scope"] == "Namespaced" and spec["subresources"] == {"status": {}}
        assert [col["name"] for col in spec["additionalPrinterColumns"]] == ["State", "Age", "Finished-TTL", "Max-Lifetime"]
        schema = spec["validation"]["openAPIV3Schema"]["properties"]
        assert schema["status"]["required"] == ["conditions", "replicaStatuses"]
        info = SCHEME.info(spec["names"]["kind"])
        assert schema["spec"]["required"] == [info.replica_specs_key]
    xdl = list(yaml.safe_load_all(open(os.path.join(str(tmp_path), "xdl.kubedl.io_xdljobs.yaml"))))[-1]
    assert "minFinishWorkRate" in xdl["spec"]["validation"]["openAPIV3Schema"]["properties"]["spec"]["properties"]


def test_crd_v1_manifests_match_registry(tmp_path):
    """The apiextensions.k8s.io/v1 form (for clusters that dropped v1beta1) stays in step too."""
    for p in crd.write_all_v1(str(tmp_path)):
        doc = list(yaml.safe_load_all(open(p)))[-1]
        committed = list(yaml.safe_load_all(open(os.path.join("config/crd/v1", os.path.basename(p)))))[-1]
        assert doc == committed, f"{p}: run python -m kubedl_b200.api.crd"
        assert doc["apiVersion"] == "apiextensions.k8s.io/v1" and "validation" not in doc["spec"]
        (v,) = doc["spec"]["versions"]
        assert v["served"] and v["storage"] and v["subresources"] == {"status": {}}
        assert [col["name"] for col in v["additionalPrinterColumns"]] == ["State", "Age", "Finished-TTL", "Max-Lifetime"]
        assert all("jsonPath" in col for col in v["additionalPrinterColumns"])

        def check(node, path="root"):      # structural: every object node has a shape or says it is free-form
            if node.get("type") == "object":
                assert "properties" in node or "additionalProperties" in node or \
                    node.get("x-kubernetes-preserve-unknown-fields"), path
            for k, child in (node.get("properties") or {}).items():
                check(child, f"{path}.{k}")
            if isinstance(node.get("additionalProperties"), dict):
                check(node["additionalProperties"], path + "[*]")
            if isinstance(node.get("items"), dict):
                check(node["items"], path + "[]")
        check(v["schema"]["openAPIV3Schema"])


def test_crd_validation_of_examples():
    import jsonschema
    for path in ("example/pytorch/pytorch_job_mnist_mpi.yaml", "example/tf/tf_job_mnist.yaml",
                 "example/xgboost/xgboostjob_v1alpha1_iris_train.yaml", "example/xdl/xdl_job_mnist.yaml"):
        for doc in yaml.safe_load_all(open("/root/reference/" + path)):
            if doc and SCHEME.is_workload(doc["kind"]):
                crd.validate(doc)
    bad = {"kind": "TFJob", "metadata": {"name": "x"}, "spec": {"cleanPodPolicy": "Sometimes", "tfReplicaSpecs": {}}}
    with pytest.raises(jsonschema.ValidationError):
        crd.validate(bad)
    with pytest.raises(jsonschema.ValidationError):
        crd.validate({"kind": "TFJob", "metadata": {"name": "x"}, "spec": {}})


def test_custom_workload_recipe():
    """API kind + ControllerInterface subclass + registry entry = a new workload."""
    from cp_helpers import Cluster, tmpl
    from kubedl_b200.controllers import registry
    from kubedl_b200.controllers.base import BaseJobReconciler
    from kubedl_b200.utils import status as st
    info = T.register_test_kind()

    class TestJobReconciler(BaseJobReconciler):
        KIND, CONTROLLER_NAME = T.TEST_JOB_KIND, "TestJobController"
        def set_cluster_spec(self, job, pod_template, rtype, index):
            for ctr in pod_template["spec"]["containers"]:
                ctr.setdefault("env", []).append({"name": "MY_INDEX", "value": index})
        def get_reconcile_orders(self): return [T.TEST_REPLICA_MASTER, T.TEST_REPLICA_WORKER]
        def is_master_role(self, replicas, rtype, index): return rtype == T.TEST_REPLICA_MASTER
        def update_job_status(self, job, replicas, job_status, restart):
            if all(rs["succeeded"] == c.replicas_of(replicas[t]) for t, rs in job_status["replicaStatuses"].items()):
                self._mark_succeeded(job, job_status, "done")
            else:
                st.update_job_conditions(job_status, c.JOB_RUNNING, st.JOB_RUNNING_REASON, "running")
    registry._register(TestJobReconciler)
    try:
        cl = Cluster()
        cl.api.install_crd(T.TEST_JOB_KIND)
        assert T.TEST_JOB_KIND in cl.recs
        job = T.new_test_job(2, name="custom", namespace="ns")
        del job["status"], job["metadata"]["uid"]
        cl.submit(job)
        assert cl.env_of(cl.api.get("Pod", "ns", "custom-worker-1"))["MY_INDEX"] == "1"
        cl.kubelet.set_all("ns", "Succeeded")
        cl.settle()
        assert cl.conditions("TestJob", "ns", "custom")[-1] == ("Succeeded", "True")
    finally:
        registry.SETUP_WITH_MANAGER_MAP.pop(T.TEST_JOB_KIND, None)
        SCHEME._by_kind.pop(T.TEST_JOB_KIND, None)


def test_deploy_manifests_are_consistent():
    """config/: the all-in-one file is what the generator emits, every kustomization points at files
    that exist, and the RBAC rules cover every resource the operator reads or writes."""
    import glob
    assert open("config/manager/all_in_one.yaml").read() == crd.all_in_one(), "run python -m kubedl_b200.api.crd"
    for path in glob.glob("config/**/*.yaml", recursive=True):
        docs = [d for d in yaml.safe_load_all(open(path)) if d]
        assert docs, path
        if os.path.basename(path) == "kustomization.yaml":
            base = os.path.dirname(path)
            for rel in (docs[0].get("resources") or []) + (docs[0].get("bases") or []) + \
                    (docs[0].get("patchesStrategicMerge") or []):
                assert os.path.exists(os.path.join(base, rel)), f"{path}: {rel} missing"
    docs = [d for d in yaml.safe_load_all(open("config/manager/all_in_one.yaml")) if d]
    allowed = set()
    for d in docs:
        if d["kind"] in ("ClusterRole", "Role"):
            for rule in d["rules"]:
                for g in rule["apiGroups"]:
                    for r in rule["resources"]:
                        allowed.add((g, r))
    needed = {("", "pods"), ("", "services"), ("", "events"), ("coordination.k8s.io", "leases"),
              ("scheduling.incubator.k8s.io", "podgroups")}
    for info in SCHEME.workload_kinds():
        needed |= {(info.group, info.plural), (info.group, info.plural + "/status")}
    assert needed <= allowed, needed - allowed
    (dep,) = [d for d in docs if d["kind"] == "Deployment"]
    pod = dep["spec"]["template"]["spec"]
    sa = [d for d in docs if d["kind"] == "ServiceAccount"][0]["metadata"]
    assert pod["serviceAccountName"] == sa["name"] and dep["metadata"]["namespace"] == sa["namespace"]
    args = pod["containers"][0]["args"]
    from kubedl_b200.cmd.manager import build_parser
    parsed = build_parser().parse_args(args)              # the Deployment's flags are flags we have
    assert parsed.in_cluster and parsed.enable_leader_election and parsed.leader_election_namespace == sa["namespace"]
    for b in [d for d in docs if d["kind"].endswith("Binding")]:
        assert b["subjects"][0]["name"] == sa["name"]
