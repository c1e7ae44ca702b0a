"""The operator against a remote API server: `KubeAPIServer` (client) <-> REST façade (server).

The reference always runs this way (client-go against kube-apiserver, reference main.go:70-75,
pkg/util/k8sutil/k8sutil.go:44-69) but never tests it.  Here the server side is the in-process
store behind `apiserver/rest.py`, which speaks the Kubernetes wire protocol.
"""
import base64
import json
import os
import threading
import time

import pytest
import yaml

from kubedl_b200.api import common as c
from kubedl_b200.api.kinds import new_job
from kubedl_b200.apiserver import kube, rest, store
from kubedl_b200.apiserver.kube import KubeAPIServer, KubeConfig

from cp_helpers import tmpl


@pytest.fixture
def remote():
    api = store.APIServer()
    srv = rest.serve(api, "127.0.0.1", 0)
    cli = KubeAPIServer(KubeConfig.from_url(srv.url), watch_timeout=2)
    yield api, srv, cli
    cli.close()
    srv.shutdown()


def wait_for(pred, timeout=10.0, what="condition"):
    t0 = time.time()
    while time.time() - t0 < timeout:
        v = pred()
        if v:
            return v
        time.sleep(0.01)
    raise AssertionError(f"timed out waiting for {what}")


def pod(name, ns="default", labels=None):
    return {"kind": "Pod", "metadata": {"name": name, "namespace": ns, "labels": labels or {}},
            "spec": {"containers": [{"name": "c", "image": "busybox"}]}}


# ---------------------------------------------------------------------------- config loading
def test_kubeconfig_token_and_inline_ca(tmp_path):
    ca = b"-----BEGIN CERTIFICATE-----\nabc\n-----END CERTIFICATE-----\n"
    doc = {"apiVersion": "v1", "kind": "Config", "current-context": "dev",
           "clusters": [{"name": "c1", "cluster": {"server": "https://10.0.0.1:6443/",
                                                   "certificate-authority-data": base64.b64encode(ca).decode()}}],
           "users": [{"name": "u1", "user": {"token": "sekret"}},
                     {"name": "u2", "user": {"client-certificate": "pki/u2.crt", "client-key": "pki/u2.key"}}],
           "contexts": [{"name": "dev", "context": {"cluster": "c1", "user": "u1", "namespace": "kubedl"}},
                        {"name": "cert", "context": {"cluster": "c1", "user": "u2"}}]}
    p = tmp_path / "config"
    p.write_text(yaml.safe_dump(doc))
    cfg = KubeConfig.from_kubeconfig(str(p))
    assert cfg.server == "https://10.0.0.1:6443" and cfg.namespace == "kubedl"
    assert cfg.headers()["Authorization"] == "Bearer sekret"
    assert open(cfg.ca_file, "rb").read() == ca
    tmp = cfg.ca_file
    cfg.cleanup()
    assert not os.path.exists(tmp)
    cert = KubeConfig.from_kubeconfig(str(p), context="cert")
    assert cert.client_cert_file == str(tmp_path / "pki" / "u2.crt")      # relative to the kubeconfig
    assert cert.client_key_file == str(tmp_path / "pki" / "u2.key")
    assert "Authorization" not in cert.headers()
    with pytest.raises(kube.ConfigError):
        KubeConfig.from_kubeconfig(str(p), context="nope")


def test_in_cluster_config_and_precedence(tmp_path):
    sa = tmp_path / "sa"
    sa.mkdir()
    (sa / "token").write_text("tok-1\n")
    (sa / "namespace").write_text("kubedl-system")
    (sa / "ca.crt").write_text("x")
    env = {"KUBERNETES_SERVICE_HOST": "10.96.0.1", "KUBERNETES_SERVICE_PORT": "443"}
    cfg = KubeConfig.in_cluster(str(sa), env)
    assert cfg.server == "https://10.96.0.1:443" and cfg.namespace == "kubedl-system"
    assert cfg.headers()["Authorization"] == "Bearer tok-1"
    (sa / "token").write_text("tok-2")                  # rotated bound token is picked up
    assert cfg.headers()["Authorization"] == "Bearer tok-2"
    with pytest.raises(kube.ConfigError):
        KubeConfig.in_cluster(str(sa), {})
    # $KUBECONFIG beats in-cluster (reference k8sutil.go:46-50); explicit flags beat both
    kc = tmp_path / "kc"
    kc.write_text(yaml.safe_dump({"current-context": "x", "clusters": [{"name": "c", "cluster": {"server": "http://h:1"}}],
                                  "contexts": [{"name": "x", "context": {"cluster": "c"}}]}))
    assert KubeConfig.load(env=dict(env, KUBECONFIG=str(kc))).server == "http://h:1"
    assert KubeConfig.load(master="127.0.0.1:9", env=dict(env, KUBECONFIG=str(kc))).server == "http://127.0.0.1:9"
    assert KubeConfig.load(master="http://m:2", kubeconfig=str(kc)).server == "http://m:2"


# ---------------------------------------------------------------------------- CRUD + errors
def test_crud_errors_and_selectors(remote):
    api, srv, cli = remote
    created = cli.create(pod("a", labels={"app": "x", "i": "0"}))
    assert created["metadata"]["uid"] and created["apiVersion"] == "v1" and created["kind"] == "Pod"
    cli.create(pod("b", labels={"app": "x", "i": "1"}))
    cli.create(pod("z", ns="other", labels={"app": "y"}))
    with pytest.raises(store.AlreadyExists):
        cli.create(pod("a"))
    with pytest.raises(store.NotFound):
        cli.get("Pod", "default", "nope")
    assert [c.name_of(p) for p in cli.list("Pod", "default")] == ["a", "b"]
    assert [c.name_of(p) for p in cli.list("Pod")] == ["a", "b", "z"]                       # all namespaces
    assert [c.name_of(p) for p in cli.list("Pod", None, {"app": "x", "i": "1"})] == ["b"]
    # optimistic concurrency
    a = cli.get("Pod", "default", "a")
    a2 = dict(a, status={"phase": "Running"})
    cli.update(a2)
    with pytest.raises(store.Conflict):
        cli.update(dict(a, status={"phase": "Failed"}))                                    # stale resourceVersion
    assert api.get("Pod", "default", "a")["status"]["phase"] == "Running"
    # merge patch of labels: set one, drop one
    out = cli.patch_labels("Pod", "default", "a", {"i": None, "role": "m"})
    assert out["metadata"]["labels"] == {"app": "x", "role": "m"}
    cli.delete("Pod", "default", "a")
    with pytest.raises(store.NotFound):
        cli.delete("Pod", "default", "a")
    with pytest.raises(store.NotFound):
        cli.list("NoSuchKind")


def test_status_subresource_and_cascade(remote):
    api, srv, cli = remote
    job = cli.create(new_job("PyTorchJob", "j", replica_specs={"Master": {"replicas": 1, "template": tmpl("pytorch")}}))
    assert job["apiVersion"] == "kubeflow.org/v1"
    # PUT on the main resource must not touch status; PUT on /status must not touch spec
    j = dict(job, status={"conditions": [{"type": "Created", "status": "True"}]})
    j = cli.update_status(j)
    assert j["status"]["conditions"][0]["type"] == "Created"
    j2 = cli.update(dict(j, status={}))
    assert j2["status"]["conditions"], "main-resource update dropped the status"
    # ownerReference garbage collection, and Orphan
    ref = {"apiVersion": job["apiVersion"], "kind": "PyTorchJob", "name": "j", "uid": job["metadata"]["uid"], "controller": True}
    p = pod("j-master-0")
    p["metadata"]["ownerReferences"] = [ref]
    cli.create(p)
    cli.delete("PyTorchJob", "default", "j")
    wait_for(lambda: not api.list("Pod"), what="dependents collected")
    job = cli.create(new_job("PyTorchJob", "k", replica_specs={}))
    p = pod("k-master-0")
    p["metadata"]["ownerReferences"] = [dict(ref, name="k", uid=job["metadata"]["uid"])]
    cli.create(p)
    cli.delete("PyTorchJob", "default", "k", cascade=False)
    assert [c.name_of(x) for x in api.list("Pod")] == ["k-master-0"]


def test_discovery_follows_crds(remote):
    api, srv, cli = remote
    cli.discovery_ttl = 0
    assert cli.crd_installed("TFJob") and cli.crd_installed("XDLJob")
    api.uninstall_crd("XDLJob")
    assert not cli.crd_installed("XDLJob")
    assert "PyTorchJob" in cli.server_resources_for_group_version("kubeflow.org/v1")
    assert cli.server_resources_for_group_version("nope.io/v1") == []
    with pytest.raises(store.NotFound):
        cli.list("XDLJob")


# ---------------------------------------------------------------------------- watch
def test_watch_replays_then_streams(remote):
    api, srv, cli = remote
    api.create(pod("pre"))
    seen = []
    stop = cli.watch("Pod", lambda ev, obj, old: seen.append((ev, c.name_of(obj), (old or {}).get("status", {}).get("phase"))))
    assert seen == [("ADDED", "pre", None)]               # replay is synchronous, like the store's
    api.create(pod("p1"))
    p = api.get("Pod", "default", "p1")
    p["status"] = {"phase": "Pending"}
    p = api.update(p)
    p["status"] = {"phase": "Running"}
    api.update(p)
    api.delete("Pod", "default", "p1")
    wait_for(lambda: len(seen) >= 5, what="watch events")
    assert seen[1:] == [("ADDED", "p1", None), ("MODIFIED", "p1", None), ("MODIFIED", "p1", "Pending"),
                        ("DELETED", "p1", None)]
    stop()
    api.create(pod("after"))
    time.sleep(0.2)
    assert len(seen) == 5
    # a second handler on the same kind shares the reflector and gets the current state
    names = []
    cli.watch("Pod", lambda ev, obj, old: names.append(c.name_of(obj)))
    assert sorted(names) == ["after", "pre"]


def test_watch_survives_stream_timeouts_and_expired_history():
    api = store.APIServer(history=16)
    srv = rest.serve(api, "127.0.0.1", 0)
    cli = KubeAPIServer(KubeConfig.from_url(srv.url), watch_timeout=1)
    try:
        state = {}

        def on(ev, obj, old):
            if ev == "DELETED":
                state.pop(c.name_of(obj), None)
            else:
                state[c.name_of(obj)] = obj["metadata"]["resourceVersion"]
        cli.watch("Pod", on)
        # events keep flowing across several 1-second watch windows
        for i in range(3):
            api.create(pod(f"p{i}"))
            time.sleep(0.6)
        wait_for(lambda: set(state) == {"p0", "p1", "p2"}, what="events across re-watches")
        # make the reflector fall off the history: cut the stream, write > history events, resume
        r = cli.reflector("Pod")
        relists = r.relists
        real_watch = r._watch_once
        gate = threading.Event()

        def held(rv):
            gate.wait(10)
            return real_watch(rv)
        r._watch_once = held
        if r._conn is not None and r._conn.sock is not None:
            r._conn.sock.shutdown(2)
        time.sleep(0.1)
        api.delete("Pod", "default", "p0")
        for i in range(40):
            api.create(pod(f"q{i}"))
        gate.set()
        wait_for(lambda: len(state) == 42 and "p0" not in state, what="re-list after 410")
        assert r.relists > relists
        assert {n: v for n, v in state.items()} == {c.name_of(o): o["metadata"]["resourceVersion"] for o in api.list("Pod")}
    finally:
        cli.close()
        srv.shutdown()


def test_events_round_trip(remote):
    api, srv, cli = remote
    job = cli.create(new_job("TFJob", "t", replica_specs={}))
    cli.record_event(job, "Normal", "SuccessfulCreatePod", "Created pod: t-worker-0", source="tf-controller")
    evs = cli.events_for(job)
    assert len(evs) == 1 and evs[0]["reason"] == "SuccessfulCreatePod"
    assert evs[0]["involvedObject"]["uid"] == job["metadata"]["uid"]
    assert evs[0]["source"]["component"] == "tf-controller"


def test_raw_wire_format(remote):
    """What kubectl / client-go see: List kinds with a resourceVersion, Status errors, watch lines."""
    import http.client
    api, srv, cli = remote
    api.create(pod("w"))
    host, port = srv.server_address[:2]
    conn = http.client.HTTPConnection(host, port, timeout=5)
    conn.request("GET", "/api/v1/namespaces/default/pods")
    d = json.loads(conn.getresponse().read())
    assert d["kind"] == "PodList" and int(d["metadata"]["resourceVersion"]) >= 1 and len(d["items"]) == 1
    conn.request("GET", "/api/v1/namespaces/default/pods/none")
    resp = conn.getresponse()
    body = json.loads(resp.read())
    assert resp.status == 404 and body["kind"] == "Status" and body["reason"] == "NotFound" and body["code"] == 404
    conn.request("GET", "/apis/kubeflow.org/v1")
    rl = json.loads(conn.getresponse().read())
    names = {r["name"] for r in rl["resources"]}
    assert {"pytorchjobs", "pytorchjobs/status", "tfjobs"} <= names and rl["groupVersion"] == "kubeflow.org/v1"
    conn.request("GET", "/apis")
    groups = {g["name"] for g in json.loads(conn.getresponse().read())["groups"]}
    assert {"kubeflow.org", "xgboostjob.kubeflow.org", "xdl.kubedl.io"} <= groups
    conn.request("GET", "/api/v1/namespaces/default/pods?watch=true&resourceVersion=0&timeoutSeconds=1")
    resp = conn.getresponse()
    line = json.loads(resp.readline())
    assert line["type"] == "ADDED" and line["object"]["metadata"]["name"] == "w"
    assert resp.read() == b""                            # stream ends after timeoutSeconds
    conn.request("GET", "/api/v1/namespaces/default/pods?watch=true&resourceVersion=999999999")
    conn.close()


# ---------------------------------------------------------------------------- operator over HTTP
def test_operator_reconciles_through_the_wire():
    """API server in one place, the operator a pure client of it (the reference's shape): a
    PyTorchJob submitted by a third party goes Created -> Running -> Succeeded, with the pods,
    services and env the in-process tests pin, all through REST + watch streams."""
    from prometheus_client import CollectorRegistry
    from kubedl_b200.cmd import manager as cmd
    from kubedl_b200.node.gpulet import FakeKubelet

    api = store.APIServer()
    srv = rest.serve(api, "127.0.0.1", 0)
    args = cmd.build_parser().parse_args(["--master", srv.url, "--workloads", "*", "--enable-leader-election=false",
                                          "--metrics-addr", "0", "--controller-metrics-addr", "0", "--serve", ""])
    reg = CollectorRegistry()
    op = cmd.build_operator(args, metrics_registry=reg)
    assert getattr(op.mgr.api, "remote", False) and op.gpulet is None
    user = KubeAPIServer(KubeConfig.from_url(srv.url))
    kubelet = FakeKubelet(api)                          # the cluster's kubelets, next to the state
    try:
        cmd.start_operator(op, args, metrics_registry=reg)
        history = []                                    # every condition list the server ever stored
        api.watch("PyTorchJob", lambda ev, obj, old: history.append(
            [x["type"] for x in obj.get("status", {}).get("conditions", [])]))
        user.create(new_job("PyTorchJob", "wire", "kubedl", replica_specs={
            "Master": {"replicas": 1, "restartPolicy": "ExitCode", "template": tmpl("pytorch")},
            "Worker": {"replicas": 2, "restartPolicy": "ExitCode", "template": tmpl("pytorch")}}))
        wait_for(lambda: len(api.list("Pod", "kubedl")) == 3, what="3 pods")
        assert [c.name_of(s) for s in api.list("Service", "kubedl")] == ["wire-master-0"]
        w1 = api.get("Pod", "kubedl", "wire-worker-1")
        env = {e["name"]: e.get("value") for e in w1["spec"]["containers"][0]["env"]}
        assert env["MASTER_ADDR"] == "wire-master-0" and env["RANK"] == "2" and env["WORLD_SIZE"] == "3"
        assert w1["metadata"]["ownerReferences"][0]["kind"] == "PyTorchJob"
        conds = lambda: [x["type"] for x in api.get("PyTorchJob", "kubedl", "wire").get("status", {}).get("conditions", [])]
        wait_for(lambda: conds() == ["Created"], what="Created")
        kubelet.set_all("kubedl", "Running")
        wait_for(lambda: conds() == ["Created", "Running"], what="Running")
        # a retryable exit code on a worker: the operator deletes the pod and recreates the index
        uid = api.get("Pod", "kubedl", "wire-worker-0")["metadata"]["uid"]
        kubelet.set_phase("kubedl", "wire-worker-0", "Failed", exit_code=137)
        wait_for(lambda: api.get("Pod", "kubedl", "wire-worker-0")["metadata"]["uid"] != uid
                 if any(c.name_of(p) == "wire-worker-0" for p in api.list("Pod", "kubedl")) else False,
                 what="worker-0 recreated")
        # Restarting is stored, then replaced by Running as soon as the master counts as running again
        wait_for(lambda: any(h and h[-1] == "Restarting" for h in history), what="Restarting recorded")
        kubelet.set_all("kubedl", "Running")
        wait_for(lambda: conds()[-1] == "Running", what="Running again")
        kubelet.set_phase("kubedl", "wire-master-0", "Succeeded", exit_code=0)
        wait_for(lambda: conds()[-1] == "Succeeded", what="Succeeded")
        job = user.get("PyTorchJob", "kubedl", "wire")
        assert job["status"]["completionTime"]
        reasons = {e["reason"] for e in user.events_for(job)}
        assert {"SuccessfulCreatePod", "SuccessfulCreateService"} <= reasons
        # client-go's rest_client_* series, by URL template rather than by object
        reg_cm = op.mgr.controller_metrics.registry
        assert reg_cm.get_sample_value("rest_client_requests_total", {
            "code": "201", "method": "POST", "host": srv.url.split("//")[1]}) >= 4     # 3 pods + 1 service (+ events)
        assert reg_cm.get_sample_value("rest_client_request_latency_seconds_count", {
            "verb": "PUT", "url": "/apis/kubeflow.org/v1/namespaces/{namespace}/pytorchjobs/{name}/status"}) >= 3
        # deleting the job through the wire collects its pods and services on the server
        user.delete("PyTorchJob", "kubedl", "wire")
        wait_for(lambda: not api.list("Pod", "kubedl") and not api.list("Service", "kubedl"), what="garbage collection")
    finally:
        op.stop()
        user.close()
        srv.shutdown()


def test_lease_leader_election(remote):
    from kubedl_b200.runtime.manager import LeaseLeaderElector
    api, srv, cli = remote
    lost = []
    a = LeaseLeaderElector(cli, "kubedl-system", identity="a", lease_duration=1.0, renew_deadline=0.6, retry_period=0.1,
                           on_lost=lambda: lost.append("a"))
    b = LeaseLeaderElector(cli, "kubedl-system", identity="b", lease_duration=1.0, renew_deadline=0.6, retry_period=0.1)
    assert a.acquire(blocking=False)
    assert not b.acquire(blocking=False)                 # a live lease is respected
    lease = api.get("Lease", "kubedl-system", "kubedl-election")
    assert lease["spec"]["holderIdentity"] == "a" and lease["spec"]["leaseTransitions"] == 0
    t0 = c.parse_time(lease["spec"]["renewTime"])
    wait_for(lambda: c.parse_time(api.get("Lease", "kubedl-system", "kubedl-election")["spec"]["renewTime"]) > t0,
             what="renewal")
    assert not b.acquire(blocking=False)
    # a releases: b takes over at once, transitions counted
    a.release()
    assert b.acquire(blocking=False)
    lease = api.get("Lease", "kubedl-system", "kubedl-election")
    assert lease["spec"]["holderIdentity"] == "b" and lease["spec"]["leaseTransitions"] == 1
    # b dies without releasing (renewals stop): after leaseDuration a new candidate wins
    b._stop.set()
    c3 = LeaseLeaderElector(cli, "kubedl-system", identity="c", lease_duration=1.0, renew_deadline=0.6, retry_period=0.1)
    assert not c3.acquire(blocking=False)
    time.sleep(1.3)
    assert c3.acquire(blocking=False)
    assert api.get("Lease", "kubedl-system", "kubedl-election")["spec"]["leaseTransitions"] == 2
    # the holder notices when somebody else owns the lease for longer than its renew deadline
    lease = api.get("Lease", "kubedl-system", "kubedl-election")
    c3.lease_duration = 30
    lease["spec"].update(holderIdentity="intruder", renewTime=c.now_precise(), leaseDurationSeconds=30)
    lost_c = []
    c3.on_lost = lambda: lost_c.append(1)
    api.update(lease)
    wait_for(lambda: lost_c, what="loss detected")
    assert not c3.is_leader
    c3.release()


def test_three_processes_run_the_reference_example(tmp_path):
    """apiserver, operator and node agent as three OS processes talking HTTP; the reference's
    PyTorch example YAML goes through `kubedlctl apply` and finishes (gloo, CPU)."""
    import socket
    import subprocess
    import sys
    from cp_helpers import REF

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    url = f"127.0.0.1:{port}"
    env = dict(os.environ, PYTHONPATH=os.getcwd() + os.pathsep + os.environ.get("PYTHONPATH", ""),
               KUBEDL_MNIST_STEPS="3", PYTHONUNBUFFERED="1")
    env.pop("KUBECONFIG", None)
    logs = {n: open(tmp_path / f"{n}.log", "w") for n in ("apiserver", "manager", "gpulet")}
    procs = []

    def spawn(name, *argv):
        p = subprocess.Popen([sys.executable, "-m", f"kubedl_b200.cmd.{name}", *argv], env=env,
                             stdout=logs[name], stderr=subprocess.STDOUT)
        procs.append(p)
        return p

    def dump():
        return "\n".join(f"--- {n}\n" + open(tmp_path / f"{n}.log").read()[-3000:] for n in logs)

    try:
        spawn("apiserver", "--serve", url)
        cli = KubeAPIServer(KubeConfig.from_url(url))
        wait_for(cli.healthy, timeout=60, what="apiserver up")
        spawn("manager", "--master", url, "--workloads", "*", "--metrics-addr", "0", "--controller-metrics-addr", "0",
              "--leader-election-namespace", "kubedl")
        spawn("gpulet", "--master", url, "--gpus", "0", "--log-dir", str(tmp_path / "pods"))
        wait_for(lambda: cli.list("Lease", "kubedl"), timeout=60, what="operator elected")
        out = subprocess.run([sys.executable, "-m", "kubedl_b200.cmd.kubedlctl", "--server", url, "apply", "-f",
                              f"{REF}/example/pytorch/pytorch_job_mnist_mpi.yaml"], env=env, capture_output=True, text=True)
        assert out.returncode == 0, out.stdout + out.stderr

        def terminal():
            for p in procs:
                assert p.poll() is None, dump()
            conds = [x["type"] for x in cli.get("PyTorchJob", "kubedl", "pytorch-dist-sendrecv-example")
                     .get("status", {}).get("conditions", [])]
            return conds if conds and conds[-1] in ("Succeeded", "Failed") else None
        try:
            conds = wait_for(terminal, timeout=180, what="job finished")
        except AssertionError:
            raise AssertionError(dump())
        assert conds == ["Created", "Running", "Succeeded"], dump()
        pods = cli.list("Pod", "kubedl")
        assert sorted(p["status"]["phase"] for p in pods) in (["Succeeded"] * 3, ["Running", "Running", "Succeeded"],
                                                             ["Running", "Succeeded", "Succeeded"])
        cli.close()
    finally:
        for p in reversed(procs):
            p.terminate()
        for p in procs:
            try:
                p.wait(timeout=10)
            except Exception:
                p.kill()
        for f in logs.values():
            f.close()


def test_install_crds_on_a_bare_server():
    """`--install-crds`: a server without our CRDs gets them (apiextensions v1), after which the
    workload gate's discovery sees the kinds (reference workload_gate.go:90-106)."""
    api = store.APIServer(install_all_crds=False)
    srv = rest.serve(api, "127.0.0.1", 0)
    cli = KubeAPIServer(KubeConfig.from_url(srv.url))
    cli.discovery_ttl = 0
    try:
        assert not cli.crd_installed("XGBoostJob")
        with pytest.raises(store.NotFound):
            cli.create(new_job("XGBoostJob", "x", replica_specs={}))
        cli.install_crd("XGBoostJob")
        cli.install_crd("XGBoostJob")                    # idempotent
        assert cli.crd_installed("XGBoostJob") and not cli.crd_installed("TFJob")
        cli.create(new_job("XGBoostJob", "x", replica_specs={}))
        crds = cli._request("GET", "/apis/apiextensions.k8s.io/v1/customresourcedefinitions")["items"]
        assert [d["metadata"]["name"] for d in crds] == ["xgboostjobs.xgboostjob.kubeflow.org"]
        v = crds[0]["spec"]["versions"][0]
        assert v["subresources"] == {"status": {}} and v["schema"]["openAPIV3Schema"]["type"] == "object"
        tmpl_schema = v["schema"]["openAPIV3Schema"]["properties"]["spec"]["properties"]["xgbReplicaSpecs"]["additionalProperties"]["properties"]["template"]
        assert tmpl_schema["x-kubernetes-preserve-unknown-fields"] is True
    finally:
        cli.close()
        srv.shutdown()


def test_unreachable_server_is_an_api_timeout_not_a_crash():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()                                           # nothing listens here
    cli = KubeAPIServer(KubeConfig.from_url(f"127.0.0.1:{port}"), request_timeout=1)
    with pytest.raises(store.Timeout):
        cli.get("Pod", "default", "x")
    assert store.is_timeout(pytest.raises(store.Timeout, cli.list, "Pod").value)
    assert not cli.healthy()
    cli.close()


def test_apiserver_snapshot_survives_a_restart(tmp_path):
    """`cmd.apiserver --state`: uids, creation timestamps, status and owner links come back, so a
    restarted operator picks its jobs up where they were and garbage collection still works."""
    from kubedl_b200.cmd import apiserver as A
    api = store.APIServer()
    job = api.create(new_job("TFJob", "t", "ns", replica_specs={"Worker": {"replicas": 1, "template": tmpl("tensorflow")}}))
    job["status"] = {"conditions": [{"type": "Created", "status": "True"}], "replicaStatuses": {}}
    api.update_status(job)
    p = pod("t-worker-0", ns="ns")
    p["metadata"]["ownerReferences"] = [{"apiVersion": "kubeflow.org/v1", "kind": "TFJob", "name": "t",
                                         "uid": job["metadata"]["uid"], "controller": True}]
    api.create(p)
    api.record_event(job, "Normal", "SuccessfulCreatePod", "x")
    path = str(tmp_path / "state.json")
    assert A.save_state(api, path) == 2                     # events are not part of the snapshot
    fresh = store.APIServer()
    assert A.load_state(fresh, path) == 2
    j2 = fresh.get("TFJob", "ns", "t")
    assert j2["metadata"]["uid"] == job["metadata"]["uid"]
    assert j2["metadata"]["creationTimestamp"] == job["metadata"]["creationTimestamp"]
    assert j2["status"]["conditions"][0]["type"] == "Created"
    fresh.delete("TFJob", "ns", "t")
    assert fresh.list("Pod", "ns") == []                    # the owner link still resolves


def test_wire_helpers():
    from kubedl_b200.apiserver.kube import _url_template, error_from_status
    from kubedl_b200.apiserver.rest import merge_patch, parse_label_selector
    assert _url_template("/api/v1/namespaces/kubedl/pods/x-worker-0?a=b") == "/api/v1/namespaces/{namespace}/pods/{name}"
    assert _url_template("/apis/kubeflow.org/v1/namespaces/ns/tfjobs/t/status") == \
        "/apis/kubeflow.org/v1/namespaces/{namespace}/tfjobs/{name}/status"
    assert _url_template("/apis/kubeflow.org/v1/tfjobs") == "/apis/kubeflow.org/v1/tfjobs"
    assert _url_template("/api/v1/namespaces/ns/pods") == "/api/v1/namespaces/{namespace}/pods"
    assert parse_label_selector("a=b, c==d,") == {"a": "b", "c": "d"} and parse_label_selector("") is None
    with pytest.raises(ValueError):
        parse_label_selector("a!=b")
    assert merge_patch({"a": {"b": 1, "c": 2}, "d": 3}, {"a": {"b": None, "e": 4}, "d": [1]}) == {"a": {"c": 2, "e": 4}, "d": [1]}
    # Status bodies -> exception classes, by reason first, by code when the body is not a Status
    assert isinstance(error_from_status(409, b'{"kind":"Status","reason":"AlreadyExists","message":"x","code":409}'), store.AlreadyExists)
    assert isinstance(error_from_status(409, b'{"kind":"Status","reason":"Conflict","message":"x","code":409}'), store.Conflict)
    assert isinstance(error_from_status(404, b"404 page not found"), store.NotFound)
    assert isinstance(error_from_status(403, b'{"reason":"Forbidden","message":"no"}'), kube.Forbidden)
    e = error_from_status(500, b"boom")
    assert isinstance(e, store.ApiError) and e.code == 500 and "boom" in str(e)


def test_operator_restart_resumes_from_api_state():
    """Kill the operator mid-job, start a new one against the same API server: it re-lists, adopts
    the pods that carry its owner reference, creates nothing twice, and drives the job to the end
    (all state lives in the API server)."""
    from prometheus_client import CollectorRegistry
    from kubedl_b200.cmd import manager as cmd
    from kubedl_b200.node.gpulet import FakeKubelet
    api = store.APIServer()
    srv = rest.serve(api, "127.0.0.1", 0)
    kubelet = FakeKubelet(api)
    flags = ["--master", srv.url, "--workloads", "*", "--enable-leader-election=false", "--metrics-addr", "0",
             "--controller-metrics-addr", "0", "--serve", ""]

    def operator():
        args = cmd.build_parser().parse_args(flags)
        reg = CollectorRegistry()
        op = cmd.build_operator(args, metrics_registry=reg)
        cmd.start_operator(op, args, metrics_registry=reg)
        return op, reg

    conds = lambda: [x["type"] for x in api.get("TFJob", "ns", "r").get("status", {}).get("conditions", [])]
    op1, _ = operator()
    try:
        api.create(new_job("TFJob", "r", "ns", replica_specs={
            "PS": {"replicas": 1, "template": tmpl("tensorflow")}, "Worker": {"replicas": 3, "template": tmpl("tensorflow")}}))
        wait_for(lambda: len(api.list("Pod", "ns")) == 4 and conds() == ["Created"], what="pods created")
        uids = {c.name_of(p): p["metadata"]["uid"] for p in api.list("Pod", "ns")}
    finally:
        op1.stop()
    # while nobody is reconciling: pods start, one worker pod disappears
    kubelet.set_all("ns", "Running")
    api.delete("Pod", "ns", "r-worker-2")
    time.sleep(0.2)
    assert conds() == ["Created"]
    creates_before = sum(1 for verb, kind, _ in api.request_log if verb == "create" and kind == "Pod")
    op2, reg2 = operator()
    try:
        wait_for(lambda: len(api.list("Pod", "ns")) == 4 and conds() == ["Created", "Running"], what="resumed")
        now = {c.name_of(p): p["metadata"]["uid"] for p in api.list("Pod", "ns")}
        assert {n: u for n, u in now.items() if n != "r-worker-2"} == {n: u for n, u in uids.items() if n != "r-worker-2"}
        assert now["r-worker-2"] != uids["r-worker-2"]                               # only the missing index was created
        assert sum(1 for verb, kind, _ in api.request_log if verb == "create" and kind == "Pod") == creates_before + 1
        assert reg2.get_sample_value("kubedl_jobs_created", {"kind": "tfjob"}) in (None, 0.0)   # Created was already recorded
        kubelet.set_all("ns", "Running")
        for i in range(3):
            kubelet.set_phase("ns", f"r-worker-{i}", "Succeeded", exit_code=0)
        wait_for(lambda: conds()[-1] == "Succeeded", what="Succeeded after restart")
    finally:
        op2.stop()
        srv.shutdown()


def test_standby_operator_takes_over(remote):
    """Two operator replicas against one API server (config/manager/in_cluster.yaml runs two): the
    standby blocks in start() until the leader goes, then reconciles what is there."""
    from prometheus_client import CollectorRegistry
    from kubedl_b200.controllers.registry import setup_with_manager
    from kubedl_b200.engine.job_controller import JobControllerConfiguration
    from kubedl_b200.runtime.manager import Manager
    api, srv, _ = remote

    def replica(identity):
        cli = KubeAPIServer(KubeConfig.from_url(srv.url), watch_timeout=2)
        mgr = Manager(api=cli, leader_election=True, leader_election_namespace="kubedl-system")
        mgr.elector.identity = identity
        mgr.elector.retry_period = 0.1
        setup_with_manager(mgr, JobControllerConfiguration(), "*", CollectorRegistry())
        return mgr

    a, b = replica("op-a"), replica("op-b")
    a.start()
    started_b = threading.Event()
    t = threading.Thread(target=lambda: (b.start(), started_b.set()), daemon=True)
    t.start()
    try:
        api.create(new_job("TFJob", "one", "ns", replica_specs={"Worker": {"replicas": 1, "template": tmpl("tensorflow")}}))
        wait_for(lambda: len(api.list("Pod", "ns")) == 1, what="leader reconciles")
        time.sleep(0.5)
        assert not started_b.is_set()                                     # the standby is waiting for the lease
        assert api.get("Lease", "kubedl-system", "kubedl-election")["spec"]["holderIdentity"] == "op-a"
        a.stop()                                                          # releases the lease
        wait_for(started_b.is_set, what="standby started")
        assert api.get("Lease", "kubedl-system", "kubedl-election")["spec"]["holderIdentity"] == "op-b"
        api.create(new_job("TFJob", "two", "ns", replica_specs={"Worker": {"replicas": 2, "template": tmpl("tensorflow")}}))
        wait_for(lambda: len(api.list("Pod", "ns")) == 3, what="new leader reconciles")
    finally:
        b.stop()
        a.stop()


def test_invalid_jobs_are_rejected_at_the_door(remote):
    """The CRD's OpenAPI schema is enforced on create and update, as kube-apiserver does: a job with a
    malformed spec is a 422 for the client, not an exception in a controller."""
    api, srv, cli = remote
    bad = new_job("TFJob", "bad", "ns", replica_specs={"Worker": {"replicas": "two", "template": tmpl("tensorflow")}})
    with pytest.raises(store.Invalid) as e:
        cli.create(bad)
    assert "replicas" in str(e.value) and "bad" in str(e.value)
    with pytest.raises(store.Invalid):
        api.create({"kind": "PyTorchJob", "metadata": {"name": "nospec", "namespace": "ns"}, "spec": {}})
    good = cli.create(new_job("TFJob", "good", "ns", replica_specs={"Worker": {"replicas": 2, "template": tmpl("tensorflow")}}))
    good["spec"]["cleanPodPolicy"] = "Sometimes"
    with pytest.raises(store.Invalid):
        cli.update(good)
    assert api.get("TFJob", "ns", "good")["spec"].get("cleanPodPolicy") != "Sometimes"
    api.validate_schema = False                          # a server without the schema takes anything
    api.create(bad)


def test_bearer_token_guards_the_api(tmp_path, capsys):
    from kubedl_b200.cmd import kubedlctl
    api = store.APIServer()
    srv = rest.serve(api, "127.0.0.1", 0, token="s3cret")
    tf = tmp_path / "token"
    tf.write_text("s3cret\n")
    anon = KubeAPIServer(KubeConfig.from_url(srv.url))
    bad = KubeAPIServer(KubeConfig.from_url(srv.url, token="nope"))
    good = KubeAPIServer(KubeConfig.load(master=srv.url, token_file=str(tf)))
    try:
        for cli in (anon, bad):
            with pytest.raises(kube.Unauthorized):
                cli.list("Pod")
            with pytest.raises(kube.Unauthorized):
                cli.create(pod("x"))
        assert anon.healthy()                                    # probes need no credentials
        good.create(pod("x"))
        assert [c.name_of(p) for p in good.list("Pod")] == ["x"] and len(api.list("Pod")) == 1
        seen = []
        good.watch("Pod", lambda ev, obj, old: seen.append(ev))   # the watch stream carries the token too
        assert seen == ["ADDED"]
        host = srv.url.split("//")[1]
        assert kubedlctl.main(["-s", host, "--token-file", str(tf), "get", "pods"]) == 0
        assert "x" in capsys.readouterr().out
    finally:
        for cli in (anon, bad, good):
            cli.close()
        srv.shutdown()


def test_shared_watchers_get_one_snapshot_private_ones_a_copy_each(remote):
    """watch(shared=True): read-only handlers of an event are handed the same object; every other handler still gets
    a copy of its own, and what it does to that copy reaches neither the other handlers nor the server.  Same contract
    on the in-process store and on the remote client."""
    api, srv, cli = remote
    for server in (api, cli):
        a, b, priv = [], [], []
        stops = [server.watch("Pod", lambda ev, obj, old: a.append(obj), replay=False, shared=True),
                 server.watch("Pod", lambda ev, obj, old: b.append(obj), replay=False, shared=True)]

        def vandal(ev, obj, old):
            priv.append(obj)
            obj["metadata"]["labels"] = {"scribbled": "yes"}
        stops.append(server.watch("Pod", vandal, replay=False))
        name = "s-" + type(server).__name__.lower()
        api.create(pod(name))
        wait_for(lambda: a and b and priv, what="watch events")
        assert a[0] is b[0] and priv[0] is not a[0]
        assert "scribbled" not in (a[0]["metadata"].get("labels") or {})
        assert "scribbled" not in (api.get("Pod", "default", name)["metadata"].get("labels") or {})
        for s in stops:
            s()


def test_merge_patch_is_one_call_status_included_and_atomic(remote):
    """merge_patch (RFC 7386) on the in-process store and through the REST client: spec, status and label deletion in
    one call; only named fields change; NotFound and a pinned resourceVersion behave as for update."""
    api, srv, cli = remote
    for server, name in ((api, "mp-local"), (cli, "mp-remote")):
        api.create(pod(name))
        o = server.merge_patch("Pod", "default", name, {"spec": {"nodeName": "n0"},
                                                        "status": {"phase": "Running", "hostIP": "127.0.0.1"},
                                                        "metadata": {"labels": {"extra": "1"}}})
        assert o["spec"]["nodeName"] == "n0" and o["status"]["phase"] == "Running" and o["metadata"]["labels"]["extra"] == "1"
        o2 = server.merge_patch("Pod", "default", name, {"status": {"phase": "Succeeded"}, "metadata": {"labels": {"extra": None}}})
        assert o2["status"] == {"phase": "Succeeded", "hostIP": "127.0.0.1"} and "extra" not in (o2["metadata"].get("labels") or {})
        assert o2["spec"]["containers"] == o["spec"]["containers"]
        with pytest.raises(store.NotFound):
            server.merge_patch("Pod", "default", "nope", {"status": {"phase": "Running"}})
        with pytest.raises(store.Conflict):
            server.merge_patch("Pod", "default", name, {"metadata": {"resourceVersion": o["metadata"]["resourceVersion"]},
                                                        "status": {"phase": "Failed"}})
        assert api.get("Pod", "default", name)["status"]["phase"] == "Succeeded"
