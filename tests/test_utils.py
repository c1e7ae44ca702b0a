"""Utility parity tests: condition state machine (reference pkg/util/status_test.go:11-101),
exit codes (pkg/util/train/train_util_test.go:19-52), workload gate
(pkg/util/workloadgate/workload_gate_test.go:8-64), naming (pkg/job_controller/util_test.go:22-32),
quota and tenancy helpers."""
import json

import pytest

from kubedl_b200.api import common as c
from kubedl_b200.engine.job_controller import gen_general_name
from kubedl_b200.utils import quota, status as st, tenancy
from kubedl_b200.utils.train import is_retryable_exit_code
from kubedl_b200.utils.workloadgate import is_workload_enable, parse_workloads_enabled


def test_condition_sequence():
    s = c.new_job_status()
    st.update_job_conditions(s, c.JOB_CREATED, st.JOB_CREATED_REASON, "created")
    assert st.is_created(s) and len(s["conditions"]) == 1
    st.update_job_conditions(s, c.JOB_RUNNING, st.JOB_RUNNING_REASON, "running")
    assert st.is_running(s) and [x["type"] for x in s["conditions"]] == ["Created", "Running"]
    # Restarting replaces Running
    st.update_job_conditions(s, c.JOB_RESTARTING, st.JOB_RESTARTING_REASON, "restarting")
    assert st.is_restarting(s) and not st.is_running(s)
    assert [x["type"] for x in s["conditions"]] == ["Created", "Restarting"]
    # and back
    st.update_job_conditions(s, c.JOB_RUNNING, st.JOB_RUNNING_REASON, "running")
    assert st.is_running(s) and not st.is_restarting(s)
    # Failed flips Running to False and freezes the list
    st.update_job_conditions(s, c.JOB_FAILED, st.JOB_FAILED_REASON, "failed")
    assert st.is_failed(s) and not st.is_running(s)
    running = st.get_condition(s, c.JOB_RUNNING)
    assert running is not None and running["status"] == "False"
    before = list(s["conditions"])
    st.update_job_conditions(s, c.JOB_RUNNING, st.JOB_RUNNING_REASON, "again")
    assert s["conditions"] == before
    assert st.last_condition_type(s) == "Failed"


def test_same_condition_is_noop_and_keeps_transition_time():
    s = c.new_job_status()
    st.update_job_conditions(s, c.JOB_RUNNING, st.JOB_RUNNING_REASON, "a")
    first = dict(s["conditions"][0])
    st.update_job_conditions(s, c.JOB_RUNNING, st.JOB_RUNNING_REASON, "b")
    assert s["conditions"][0] == first                      # same type/status/reason -> untouched
    st.update_job_conditions(s, c.JOB_RUNNING, "OtherReason", "c")
    assert s["conditions"][0]["reason"] == "OtherReason"
    assert s["conditions"][0]["lastTransitionTime"] == first["lastTransitionTime"]


def test_succeeded_flips_running_false():
    s = c.new_job_status()
    st.update_job_conditions(s, c.JOB_RUNNING, st.JOB_RUNNING_REASON, "r")
    st.update_job_conditions(s, c.JOB_SUCCEEDED, st.JOB_SUCCEEDED_REASON, "s")
    assert st.is_succeeded(s) and not st.is_running(s)


@pytest.mark.parametrize("code,retry", [(1, False), (2, False), (3, False), (126, False), (127, False),
                                        (128, False), (139, False), (130, True), (137, True), (143, True),
                                        (138, True), (0, False), (255, False)])
def test_exit_codes(code, retry):
    assert is_retryable_exit_code(code) is retry


@pytest.mark.parametrize("spec,enables,enable_all", [
    ("", {}, False), ("*", {}, True), ("*,foo", {"foo": True}, True), ("foo,*", {"foo": True}, True),
    ("foo,a", {"foo": True, "a": True}, False), ("foo,-a", {"foo": True, "a": False}, False),
    ("-foo,a", {"foo": False, "a": True}, False), ("foo,-*", {"foo": True}, False)])
def test_parse_workloads(spec, enables, enable_all):
    assert parse_workloads_enabled(spec) == (enables, enable_all)


def test_workload_gate(monkeypatch):
    monkeypatch.delenv("WORKLOADS_ENABLE", raising=False)
    monkeypatch.delenv("KUBEDL_CI", raising=False)
    assert is_workload_enable("TFJob", "auto", crd_installed=lambda kind: kind == "TFJob")
    assert not is_workload_enable("XDLJob", "auto", crd_installed=lambda kind: kind == "TFJob")
    assert is_workload_enable("XDLJob", "*")
    assert not is_workload_enable("XDLJob", "*,-XDLJob")
    assert is_workload_enable("TFJob", "TFJob,PyTorchJob") and not is_workload_enable("XDLJob", "TFJob")
    monkeypatch.setenv("WORKLOADS_ENABLE", "XDLJob")          # env overrides the flag
    assert is_workload_enable("XDLJob", "TFJob") and not is_workload_enable("TFJob", "TFJob")
    monkeypatch.setenv("WORKLOADS_ENABLE", "auto")
    monkeypatch.setenv("KUBEDL_CI", "true")                   # CI: no discovery client -> enabled
    assert is_workload_enable("XDLJob", "auto", crd_installed=lambda kind: False)


def test_gen_general_name():
    assert gen_general_name("job", "worker", "0") == "job-worker-0"
    assert gen_general_name("ns/job", "ps", "3") == "ns-job-ps-3"


def test_quota_sum_and_max():
    pod = {"containers": [{"resources": {"requests": {"cpu": "1024m", "memory": "1Gi"}, "limits": {"cpu": "2"}}},
                          {"resources": {"requests": {"cpu": "1", "memory": "512Mi"}}}],
           "initContainers": [{"resources": {"requests": {"cpu": "4", "memory": "100Mi"}}}]}
    r = quota.pod_resources(pod)
    assert abs(r["requests"]["cpu"] - 4.0) < 1e-9                  # max(sum containers 2.024, init 4)
    assert r["requests"]["memory"] == 2**30 + 512 * 2**20           # sum wins over init
    assert r["limits"]["cpu"] == 2.0
    wire = quota.resources_to_wire(r)
    assert wire["requests"]["cpu"] == "4" and wire["requests"]["memory"] == "1536Mi"


def test_tenancy_annotation():
    obj = {"metadata": {"annotations": {c.ANNOTATION_TENANCY_INFO: json.dumps(
        {"tenant": "t", "user": "u", "idc": "i", "region": "r"})}}}
    t = tenancy.get_tenancy(obj)
    assert (t.tenant, t.user, t.idc, t.region) == ("t", "u", "i", "r")
    assert tenancy.get_tenancy({"metadata": {}}) is None
    with pytest.raises(ValueError):
        tenancy.get_tenancy({"metadata": {"annotations": {c.ANNOTATION_TENANCY_INFO: "{bad"}}})


def test_stem_weight_pad_unpad_roundtrip():
    """The stem filter's patch-column layout (7 filter rows x 24 columns, 21 used) is a pure re-indexing."""
    import torch
    from kubedl_b200.ops import nn as K
    w = torch.randn(64, 7, 7, 3)
    wp = K.stem_weight_padded(w)
    assert wp.shape == (64, K.STEM_KP) and K.STEM_KP == 168
    assert float(wp.view(64, 7, K.STEM_KW)[:, :, 21:].abs().max()) == 0.0
    assert torch.equal(K.stem_weight_unpad(wp).view(64, 7, 7, 3), w)
    # column of (kh, kw, c) is kh*24 + kw*3 + c
    assert float(wp[5, 2 * 24 + 4 * 3 + 1]) == float(w[5, 2, 4, 1])


def test_worker_exit_code_classification():
    """What a worker dies of decides whether restartPolicy ExitCode brings it back
    (reference pkg/util/train/train_util.go:18-53)."""
    from kubedl_b200.utils.train import is_retryable_exit_code
    from kubedl_b200.worker.exitcodes import classify, run
    retry = [RuntimeError("CUDA error: unspecified launch failure"), RuntimeError("NCCL error in: ... unhandled system error"),
             ConnectionResetError("peer"), TimeoutError(), RuntimeError("Socket Timeout"),
             RuntimeError("[c10d] the client socket has timed out after 600s while trying to connect to the store"),
             RuntimeError("all-reduce watchdog: peer did not arrive"), KeyboardInterrupt(), MemoryError(),
             type("DistStoreError", (RuntimeError,), {})("wait timeout after 15000ms, keys: /default_pg/0//cpu//0/0"),
             RuntimeError("[gloo/transport/tcp/pair.cc:547] Connection closed by peer [127.0.0.1]:59032"),
             RuntimeError("[/pytorch/third_party/gloo/gloo/transport/tcp/pair.cc:311] Connect timeout [none]")]
    for e in retry:
        assert is_retryable_exit_code(classify(e)), e
    perm = [ValueError("batch must be a multiple of 8"), ImportError("no module"), AssertionError("shape"),
            RuntimeError("CUDA out of memory. Tried to allocate 2.00 GiB"), KeyError("RANK")]
    for e in perm:
        assert not is_retryable_exit_code(classify(e)), e
    assert classify(SystemExit(0)) == 0 and classify(SystemExit(3)) == 3 and classify(SystemExit(None)) == 0
    assert run(lambda: None) == 0 and run(lambda: 7) == 7

    def boom():
        raise RuntimeError("CUDA error: an illegal memory access was encountered")
    assert run(boom) == 138
