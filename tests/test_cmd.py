"""Operator entry, flags (reference main.go:54-60, docs/startup_flags.md) and the kubedlctl client
against the REST façade."""
import json
import socket
import time

import pytest
from prometheus_client import CollectorRegistry

from cp_helpers import REF
from kubedl_b200.cmd import kubedlctl, manager


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def test_flag_defaults_match_reference():
    a = manager.build_parser().parse_args([])
    assert a.controller_metrics_addr == ":8080" and a.metrics_addr == 8443
    assert a.enable_leader_election is True and a.gang_scheduler_name == "" and a.max_reconciles == 1
    assert a.workloads == "auto" and a.region == "" and a.event_storage == "" and a.object_storage == ""
    a = manager.build_parser().parse_args(["--enable-leader-election=false", "--max-reconciles", "0"])
    assert a.enable_leader_election is False


def test_operator_wiring_and_kubedlctl(capsys, tmp_path):
    port = _free_port()
    args = manager.build_parser().parse_args(["--enable-leader-election=false", "--serve", f"127.0.0.1:{port}",
                                              "--metrics-addr", "0", "--controller-metrics-addr", "0",
                                              "--workloads", "*", "--no-node-agent", "--max-reconciles", "0", "--gang-scheduler-name", "kube-batch",
                                              "--object-storage", "mysql", "--region", "box"])
    reg = CollectorRegistry()
    op = manager.build_operator(args, metrics_registry=reg)
    assert manager.CTRL_CONFIG.max_concurrent_reconciles == 1 and manager.CTRL_CONFIG.enable_gang_scheduling
    assert set(op.reconcilers) == {"PyTorchJob", "TFJob", "XGBoostJob", "XDLJob"} and "pod" in op.persisters
    manager.start_operator(op, args, metrics_registry=reg)
    try:
        srv = f"127.0.0.1:{port}"
        assert kubedlctl.main(["-s", srv, "apply", "-f", f"{REF}/example/xdl/xdl_job_mnist.yaml"]) == 0
        out = capsys.readouterr().out
        # the ZooKeeper Service + Deployment in the same file are stored (nothing runs them here)
        assert "xdljob/xdl-mnist-example applied" in out and "deployment/zk-0 applied" in out
        for _ in range(100):
            kubedlctl.main(["-s", srv, "get", "pods", "-n", "kubedl"])
            if "xdl-mnist-example-worker-1" in capsys.readouterr().out:
                break
            time.sleep(0.05)
        else:
            raise AssertionError("pods never appeared")
        assert kubedlctl.main(["-s", srv, "get", "xdljobs", "-n", "kubedl"]) == 0
        table = capsys.readouterr().out
        assert "NAME" in table and "STATE" in table and "FINISHED-TTL" in table and "3600" in table
        kubedlctl.main(["-s", srv, "get", "xdljob", "xdl-mnist-example", "-n", "kubedl", "-o", "json"])
        obj = json.loads(capsys.readouterr().out)
        assert obj["kind"] == "XDLJob" and obj["spec"]["ttlSecondsAfterFinished"] == 3600
        assert kubedlctl.main(["-s", srv, "describe", "xdljob", "xdl-mnist-example", "-n", "kubedl"]) == 0
        described = capsys.readouterr().out
        assert "SuccessfulCreatePod" in described
        assert "Pods:" in described and "xdl-mnist-example-worker-1" in described and "RESTARTS" in described
        assert kubedlctl.main(["-s", srv, "get", "podgroups", "-n", "kubedl"]) == 0
        assert "xdl-mnist-example" in capsys.readouterr().out
        assert kubedlctl.main(["-s", srv, "delete", "xdljob", "xdl-mnist-example", "-n", "kubedl"]) == 0
        with pytest.raises(SystemExit):
            kubedlctl.main(["-s", srv, "get", "xdljob", "xdl-mnist-example", "-n", "kubedl"])
        with pytest.raises(SystemExit):
            kubedlctl.resolve_kind("nonsense")
    finally:
        op.stop()
        manager.CTRL_CONFIG.enable_gang_scheduling = False
        manager.CTRL_CONFIG.gang_scheduler_name = ""


def test_leader_election_is_exclusive(tmp_path):
    from kubedl_b200.runtime.manager import LeaderElector
    lock = str(tmp_path / "leader.lock")
    a, b = LeaderElector(lock, "a"), LeaderElector(lock, "b")
    assert a.acquire(blocking=False) is True
    assert b.acquire(blocking=False) is False
    a.release()
    assert b.acquire(blocking=False) is True
    b.release()


def test_controller_runtime_metrics_endpoint():
    """`--controller-metrics-addr` serves what controller-runtime serves for the reference
    (main.go:70-72): reconcile and work-queue series per controller, by the upstream names."""
    import urllib.request
    from kubedl_b200.api.kinds import new_job
    from kubedl_b200.node.gpulet import FakeKubelet
    from cp_helpers import tmpl
    args = manager.build_parser().parse_args(["--enable-leader-election=false", "--serve", "", "--metrics-addr", "0",
                                              "--controller-metrics-addr", "127.0.0.1:0", "--workloads", "*", "--no-node-agent"])
    op = manager.build_operator(args, metrics_registry=CollectorRegistry())
    manager.start_operator(op, args)
    try:
        (srv,) = op.servers
        url = "http://%s:%d/metrics" % srv.server_address[:2]
        api = op.mgr.api
        api.create(new_job("TFJob", "m", "ns", replica_specs={"Worker": {"replicas": 2, "template": tmpl("tensorflow")}}))
        kubelet = FakeKubelet(api)
        for _ in range(200):
            if len(api.list("Pod", "ns")) == 2:
                break
            time.sleep(0.01)
        kubelet.set_all("ns", "Running")
        kubelet.set_all("ns", "Succeeded", exit_code=0)
        for _ in range(300):
            conds = [x["type"] for x in api.get("TFJob", "ns", "m").get("status", {}).get("conditions", [])]
            if conds and conds[-1] == "Succeeded":
                break
            time.sleep(0.01)
        assert conds[-1] == "Succeeded"
        text = urllib.request.urlopen(url, timeout=5).read().decode()
        samples = {}
        for line in text.splitlines():
            if line and not line.startswith("#"):
                name, _, value = line.rpartition(" ")
                samples[name] = float(value)
        ok = samples['controller_runtime_reconcile_total{controller="TFJobController",result="success"}']
        assert ok >= 3
        assert samples['controller_runtime_reconcile_time_seconds_count{controller="TFJobController"}'] >= ok
        assert samples['workqueue_adds_total{name="TFJobController"}'] >= ok
        assert samples['workqueue_depth{name="TFJobController"}'] == 0
        assert samples['workqueue_work_duration_seconds_count{name="TFJobController"}'] >= ok
        assert samples['workqueue_queue_duration_seconds_count{name="TFJobController"}'] >= ok
        assert 'workqueue_longest_running_processor_seconds{name="TFJobController"}' in samples
        assert 'workqueue_depth{name="PytorchJobController"}' in samples       # every controller has its queue
        base = url.rsplit("/", 1)[0]
        trace = json.loads(urllib.request.urlopen(base + "/debug/trace", timeout=5).read())
        assert trace["TFJobController"] and trace["TFJobController"][-1]["key"] == "ns/m"
        assert all(sp["error"] is None and sp["seconds"] >= 0 for sp in trace["TFJobController"])
        queues = json.loads(urllib.request.urlopen(base + "/debug/queues", timeout=5).read())
        assert queues["TFJobController"]["depth"] == 0 and queues["TFJobController"]["reconciles"] >= ok
        # an error is counted as such and as a retry
        from kubedl_b200.apiserver.store import Invalid
        api.create(new_job("TFJob", "bad", "ns", replica_specs={"Worker": {"replicas": 1, "template": tmpl("tensorflow")}}))
        time.sleep(0.1)
        api.fail_next[("create", "Pod")] = Invalid("nope")
        api.delete("Pod", "ns", "bad-worker-0")
        for _ in range(300):
            text = urllib.request.urlopen(url, timeout=5).read().decode()
            if 'controller_runtime_reconcile_errors_total{controller="TFJobController"} 1.0' in text:
                break
            time.sleep(0.01)
        assert 'controller_runtime_reconcile_errors_total{controller="TFJobController"} 1.0' in text
        assert 'workqueue_retries_total{name="TFJobController"} 1.0' in text
    finally:
        op.stop()


def test_bind_address_parsing():
    from kubedl_b200.plugins.controller_metrics import parse_bind_address
    assert parse_bind_address(":8080") == ("0.0.0.0", 8080)
    assert parse_bind_address("127.0.0.1:9090") == ("127.0.0.1", 9090)
    assert parse_bind_address("0") is None and parse_bind_address("") is None


def test_kubedlctl_with_a_kubeconfig(capsys, tmp_path):
    """`kubedlctl --kubeconfig`: the same verbs through the Kubernetes client (here against the façade,
    which speaks the same protocol as kube-apiserver)."""
    import yaml
    from kubedl_b200.apiserver import rest, store
    api = store.APIServer()
    srv = rest.serve(api, "127.0.0.1", 0, log_reader=lambda ns, name: f"log of {ns}/{name}\n")
    kc = tmp_path / "config"
    kc.write_text(yaml.safe_dump({"current-context": "box", "clusters": [{"name": "b", "cluster": {"server": srv.url}}],
                                  "contexts": [{"name": "box", "context": {"cluster": "b", "namespace": "kubedl"}}]}))
    base = ["--kubeconfig", str(kc)]
    try:
        assert kubedlctl.main(base + ["apply", "-f", f"{REF}/example/pytorch/pytorch_job_mnist_mpi.yaml"]) == 0
        assert "pytorchjob/pytorch-dist-sendrecv-example applied" in capsys.readouterr().out
        assert api.get("PyTorchJob", "kubedl", "pytorch-dist-sendrecv-example")
        assert kubedlctl.main(base + ["get", "pytorchjobs"]) == 0                 # namespace from the context
        assert "pytorch-dist-sendrecv-example" in capsys.readouterr().out
        kubedlctl.main(base + ["get", "pytorchjob", "pytorch-dist-sendrecv-example", "-o", "json"])
        assert json.loads(capsys.readouterr().out)["kind"] == "PyTorchJob"
        api.create({"kind": "Pod", "metadata": {"name": "p0", "namespace": "kubedl"}, "spec": {"containers": []}})
        assert kubedlctl.main(base + ["logs", "p0"]) == 0
        assert capsys.readouterr().out == "log of kubedl/p0\n"
        assert kubedlctl.main(base + ["describe", "pytorchjob", "pytorch-dist-sendrecv-example"]) == 0
        capsys.readouterr()
        assert kubedlctl.main(base + ["delete", "pytorchjob", "pytorch-dist-sendrecv-example"]) == 0
        with pytest.raises(SystemExit) as e:
            kubedlctl.main(base + ["get", "pytorchjob", "pytorch-dist-sendrecv-example"])
        assert "NotFound" in str(e.value)
    finally:
        srv.shutdown()


def test_apiserver_refuses_an_empty_token_file(tmp_path, capsys):
    """`--token-file` naming a file with no token in it must not start an unauthenticated server."""
    from kubedl_b200.cmd import apiserver
    tf = tmp_path / "token"
    tf.write_text("  \n")
    assert apiserver.main(["--serve", "127.0.0.1:0", "--token-file", str(tf)]) == 2
    assert "holds no token" in capsys.readouterr().err


def test_kubedlctl_get_watch_prints_a_row_per_change(capsys):
    """`kubedlctl get -w`: the table, then one row for every change that arrives on the API server's watch stream
    (what `kubectl get -w` does against the reference's CRDs with their printer columns)."""
    import threading
    from kubedl_b200.bench.launch_delay import _job
    port = _free_port()
    args = manager.build_parser().parse_args(["--enable-leader-election=false", "--serve", f"127.0.0.1:{port}",
                                              "--metrics-addr", "0", "--controller-metrics-addr", "0",
                                              "--workloads", "*", "--no-node-agent"])
    reg = CollectorRegistry()
    op = manager.build_operator(args, metrics_registry=reg)
    manager.start_operator(op, args, metrics_registry=reg)
    try:
        srv = f"127.0.0.1:{port}"
        op.mgr.api.create(_job("PyTorchJob", "first", 2, 0.1))

        def later():
            time.sleep(0.4)
            op.mgr.api.create(_job("PyTorchJob", "second", 2, 0.1))
            time.sleep(0.3)
            op.mgr.api.delete("PyTorchJob", "bench", "first")
        t = threading.Thread(target=later)
        t.start()
        assert kubedlctl.main(["-s", srv, "get", "pytorchjobs", "-n", "bench", "-w", "--watch-seconds", "1.5"]) == 0
        t.join()
        out = capsys.readouterr().out.splitlines()
        assert out[0].split()[:3] == ["NAME", "STATE", "AGE"] and out[1].startswith("first")
        rest = out[2:]
        assert any(l.startswith("second") and "Created" in l for l in rest), out
        assert any(l.startswith("first") and "<deleted>" in l for l in rest), out
        # following one object only
        assert kubedlctl.main(["-s", srv, "get", "pytorchjob", "second", "-n", "bench", "-w", "--watch-seconds", "0.3"]) == 0
        assert all(l.startswith(("NAME", "second")) for l in capsys.readouterr().out.splitlines())
    finally:
        op.stop()


def test_kubedlctl_logs_follow_prints_until_the_pod_finishes(capsys):
    """`kubedlctl logs -f`: the log so far, then what the worker appends, until its pod is Succeeded."""
    import sys
    from kubedl_b200.api import kinds as k
    port = _free_port()
    args = manager.build_parser().parse_args(["--enable-leader-election=false", "--serve", f"127.0.0.1:{port}",
                                              "--metrics-addr", "0", "--controller-metrics-addr", "0",
                                              "--workloads", "*", "--gpus", "0"])
    reg = CollectorRegistry()
    op = manager.build_operator(args, metrics_registry=reg)
    manager.start_operator(op, args, metrics_registry=reg)
    try:
        srv = f"127.0.0.1:{port}"
        code = "import time\nfor i in range(4):\n    print('tick', i, flush=True); time.sleep(0.25)"
        t = {"spec": {"containers": [{"name": "xdl", "image": "python", "command": [sys.executable, "-c", code]}]}}
        import uuid
        job = "talker-" + uuid.uuid4().hex[:6]          # pod logs are files named after the pod and appended to
        op.mgr.api.create(k.new_job("XDLJob", job, "bench", {"Worker": {"replicas": 1, "restartPolicy": "Never", "template": t}}))
        for _ in range(200):
            if op.mgr.api.list("Pod", "bench"):
                break
            time.sleep(0.02)
        assert kubedlctl.main(["-s", srv, "logs", f"{job}-worker-0", "-n", "bench", "-f", "--follow-seconds", "20"]) == 0
        out = capsys.readouterr().out
        assert [l for l in out.splitlines() if l.startswith("tick")] == [f"tick {i}" for i in range(4)], out
    finally:
        op.stop()


def test_every_flag_is_documented():
    """docs/startup_flags.md (the reference's docs/startup_flags.md:4-9, extended) names every flag of the three commands."""
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    doc = open(os.path.join(root, "docs", "startup_flags.md")).read()
    flags = {o for a in manager.build_parser()._actions for o in a.option_strings if o.startswith("--")} - {"--help"}
    for mod in ("gpulet", "apiserver"):
        flags |= set(re.findall(r'add_argument\("(--[a-z\-]+)"', open(os.path.join(root, "kubedl_b200", "cmd", mod + ".py")).read()))
    assert sorted(f for f in flags if f not in doc) == []
