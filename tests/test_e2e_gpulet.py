"""End to end on CPU (BASELINE.json config #1): the reference's PyTorchJob example YAML, unchanged,
through the controllers and the node agent — real processes, real gloo rendezvous wired only by
the injected env — until the job reports Succeeded; plus the gpulet helpers."""
import os
import sys
import time

import pytest
from prometheus_client import CollectorRegistry

from cp_helpers import load_example, tmpl
from kubedl_b200.api import common as c
from kubedl_b200.api import kinds as k
from kubedl_b200.controllers.registry import setup_with_manager
from kubedl_b200.engine.job_controller import JobControllerConfiguration
from kubedl_b200.node.gpulet import Gpulet, ImageRegistry, expand_vars
from kubedl_b200.plugins import gang_schedule
from kubedl_b200.runtime.manager import Manager


def _wait(fn, timeout=120.0, every=0.05):
    t0 = time.time()
    while time.time() - t0 < timeout:
        v = fn()
        if v:
            return v
        time.sleep(every)
    raise AssertionError("timed out")


def _cluster(num_gpus=0, gang=""):
    mgr = Manager()
    gang_schedule.register_gang_schedulers(mgr)
    cfg = JobControllerConfiguration(enable_gang_scheduling=bool(gang), gang_scheduler_name=gang)
    recs = setup_with_manager(mgr, cfg, "*", CollectorRegistry())
    let = Gpulet(mgr.api, num_gpus=num_gpus, log_dir=f"/tmp/kubedl-b200/test-pods-{os.getpid()}")
    mgr.start()
    let.start()
    return mgr, recs, let


def _terminal(mgr, kind, ns, name):
    conds = [x["type"] for x in mgr.api.get(kind, ns, name).get("status", {}).get("conditions", [])]
    return conds if ("Succeeded" in conds or "Failed" in conds) else None


def test_expand_vars():
    env = {"TASK_NAME": "worker", "TASK_INDEX": "1"}
    assert expand_vars("--task_name $(TASK_NAME) --i $(TASK_INDEX) $(NOPE) $$(TASK_NAME)", env) == \
        "--task_name worker --i 1 $(NOPE) $(TASK_NAME)"


def test_image_registry_resolution():
    reg = ImageRegistry()
    argv = reg.resolve({"image": "kubedl/pytorch-dist-example"}, {})
    assert argv[-2:] == ["-m", "kubedl_b200.worker.mnist_ddp"]
    argv = reg.resolve({"image": "busybox", "command": ["sh", "-c"], "args": ["echo $(X)"]}, {"X": "hi"})
    assert argv == ["sh", "-c", "echo hi"]
    with pytest.raises(RuntimeError):
        reg.resolve({"image": "unknown/image"}, {})


def test_pytorch_example_end_to_end_gloo():
    mgr, recs, let = _cluster()
    try:
        (job,) = load_example("example/pytorch/pytorch_job_mnist_mpi.yaml")
        for spec in job["spec"]["pytorchReplicaSpecs"].values():          # keep the run short
            spec["template"]["spec"]["containers"][0]["env"] = [{"name": "KUBEDL_MNIST_STEPS", "value": "5"}]
        mgr.api.create(job)
        conds = _wait(lambda: _terminal(mgr, "PyTorchJob", "kubedl", "pytorch-dist-sendrecv-example"))
        assert conds == ["Created", "Running", "Succeeded"], (conds, let.pod_log("kubedl", "pytorch-dist-sendrecv-example-master-0"))
        st = mgr.api.get("PyTorchJob", "kubedl", "pytorch-dist-sendrecv-example")["status"]
        assert st["replicaStatuses"]["Master"]["succeeded"] == 1
        log = let.pod_log("kubedl", "pytorch-dist-sendrecv-example-worker-1")
        assert "rank 2/3" in log
        m = recs["PyTorchJob"].metrics
        assert len(m.launch_delays["first_pod"]) == 1 and len(m.launch_delays["all_pods"]) == 1
        assert 0 <= m.launch_delays["first_pod"][0][1] <= m.launch_delays["all_pods"][0][1] < 30
    finally:
        let.stop()
        mgr.stop()


def test_exit_code_restart_end_to_end():
    """Worker exits with the user-retryable code 138 once -> pod recreated with the same rank ->
    job still succeeds (reference pod.go:282-304, train_util.go:18-53)."""
    mgr, recs, let = _cluster()
    try:
        marker = f"/tmp/kubedl-b200/restarted-{os.getpid()}"
        if os.path.exists(marker):
            os.remove(marker)
        script = (f"import os,sys\n"
                  f"if os.environ['KUBEDL_REPLICA_TYPE']=='worker' and not os.path.exists('{marker}'):\n"
                  f"    open('{marker}','w').close(); sys.exit(138)\n"
                  f"sys.exit(0)\n")
        t = {"spec": {"containers": [{"name": "pytorch", "image": "python", "command": ["python", "-c", script]}]}}
        job = k.new_job("PyTorchJob", "flaky", "ns", {"Master": {"restartPolicy": "ExitCode", "template": t},
                                                       "Worker": {"restartPolicy": "ExitCode", "template": t}})
        # keep the master alive until the worker has been through its restart
        job["spec"]["pytorchReplicaSpecs"]["Master"]["template"] = {"spec": {"containers": [{
            "name": "pytorch", "image": "python",
            "command": ["python", "-c", f"import os,time\nwhile not os.path.exists('{marker}'): time.sleep(0.05)\ntime.sleep(1.0)"]}]}}
        mgr.api.create(job)
        conds = _wait(lambda: _terminal(mgr, "PyTorchJob", "ns", "flaky"))
        assert conds[-1] == "Succeeded", conds
        assert "JobRestarting" in recs["PyTorchJob"].recorder.reasons()
        assert "ExitedWithCode" in recs["PyTorchJob"].recorder.reasons()
    finally:
        let.stop()
        mgr.stop()


def test_gpu_gang_admission_is_all_or_nothing():
    """Two 2-GPU gangs on a 3-GPU node: the second gang must wait for the first to finish."""
    mgr, recs, let = _cluster(num_gpus=3, gang="gpu-gang")
    try:
        def job(name, seconds):
            t = {"spec": {"containers": [{"name": "pytorch", "image": "python",
                                          "command": ["python", "-c", f"import time; time.sleep({seconds})"],
                                          "resources": {"limits": {"nvidia.com/gpu": 1}}}]}}
            return k.new_job("PyTorchJob", name, "ns", {"Master": {"restartPolicy": "Never", "template": t},
                                                         "Worker": {"restartPolicy": "Never", "template": t}},
                             cleanPodPolicy="All")
        mgr.api.create(job("a", 1.5))
        _wait(lambda: len([p for p in mgr.api.list("Pod", "ns") if p.get("status", {}).get("phase") == "Running"]) == 2)
        mgr.api.create(job("b", 0.2))
        time.sleep(0.5)
        b_pods = [p for p in mgr.api.list("Pod", "ns", {"job-name": "b"})]
        assert len(b_pods) == 2 and all(not p.get("status", {}).get("phase") or p["status"]["phase"] == "Pending"
                                        for p in b_pods)                 # 1 free GPU < 2 needed: nobody starts
        assert mgr.api.get("PodGroup", "ns", "b")["spec"]["minResources"] == {"nvidia.com/gpu": 2}
        _wait(lambda: _terminal(mgr, "PyTorchJob", "ns", "a"))
        assert _wait(lambda: _terminal(mgr, "PyTorchJob", "ns", "b"))[-1] == "Succeeded"
        # a GPU returns to the pool when its pod's process is really gone (a still-exiting worker deleted by
        # cleanPodPolicy All is reaped on the agent's next pass), not at the moment of the delete
        _wait(lambda: set(let.free_gpus) == {0, 1, 2}, timeout=15)
        assert not let._terminating and not let._gpu_of
    finally:
        let.stop()
        mgr.stop()


def test_rank_failure_recovers_through_exit_codes(tmp_path):
    """A rank dies mid-run with the retryable 138.  Its peers' collectives fail, `worker/exitcodes.py`
    turns those failures into 138 as well, the ExitCode policy recreates every pod that left that
    way (same names, same ranks), the job rendezvouses again and finishes."""
    mgr, recs, let = _cluster()
    try:
        (job,) = load_example("example/pytorch/pytorch_job_mnist_mpi.yaml")
        job["metadata"]["name"] = "recover"
        for spec in job["spec"]["pytorchReplicaSpecs"].values():
            spec["template"]["spec"]["containers"][0]["env"] = [
                {"name": "KUBEDL_MNIST_STEPS", "value": "6"}, {"name": "KUBEDL_FAIL_WITH", "value": "138"},
                {"name": "KUBEDL_FAIL_ONCE", "value": str(tmp_path / "failed-once")},
                # a recreated rank that met the dying incarnation's store gives up quickly and retries
                {"name": "KUBEDL_RENDEZVOUS_TIMEOUT", "value": "8"}]
        seen = []
        mgr.api.watch("PyTorchJob", lambda ev, obj, old: seen.append(
            [x["type"] for x in obj.get("status", {}).get("conditions", [])][-1:]))
        mgr.api.create(job)
        conds = _wait(lambda: _terminal(mgr, "PyTorchJob", "kubedl", "recover"), timeout=180)
        logs = {p: let.pod_log("kubedl", f"recover-{p}") for p in ("master-0", "worker-0", "worker-1")}
        assert conds[-1] == "Succeeded", (conds, logs)
        assert ["Restarting"] in seen
        assert "injected failure with exit code 138" in logs["worker-1"]
        assert "exiting with 138 (retryable" in logs["master-0"] + logs["worker-0"]     # a survivor left retryably
        assert logs["worker-1"].count("rank 2/3: 6 steps") == 1                           # a later incarnation finished
    finally:
        let.stop()
        mgr.stop()


def test_xgboost_train_then_predict_jobs(tmp_path, monkeypatch):
    """The reference's XGBoost example has two job types, Train and Predict, linked by --model_path
    (example/xgboost/xgboostjob_v1alpha1_iris_train.yaml:21-28).  Train over 3 ranks (histograms
    summed with gloo, every rank grows identical trees), rank 0 writes the model; a second
    XGBoostJob loads it and scores the same shards: same loss."""
    import json as _json
    monkeypatch.setenv("KUBEDL_MODEL_DIR", str(tmp_path / "models"))
    mgr, recs, let = _cluster()
    try:
        (job,) = load_example("example/xgboost/xgboostjob_v1alpha1_iris_train.yaml")
        ns, name = job["metadata"].get("namespace", "default"), job["metadata"]["name"]
        for spec in job["spec"]["xgbReplicaSpecs"].values():
            ctr = spec["template"]["spec"]["containers"][0]
            ctr["env"] = [{"name": "KUBEDL_XGB_ROWS", "value": "4000"}]
            ctr["args"] = [a for a in ctr["args"] if not a.startswith("--model_path")] + ["--model_path=autoAI/xgb-opt/2"]
        mgr.api.create(job)
        conds = _wait(lambda: _terminal(mgr, "XGBoostJob", ns, name), timeout=180)
        logs = [let.pod_log(ns, f"{name}-{p}") for p in ("master-0", "worker-0", "worker-1")]
        assert conds[-1] == "Succeeded", logs
        res = [_json.loads([l for l in lg.splitlines() if l.startswith("{")][-1]) for lg in logs]
        assert [r["rank"] for r in res] == [0, 1, 2] and all(r["world"] == 3 and r["trees"] == 3 for r in res)
        assert "written to" in logs[0] and "kept on local disk" in logs[0]          # oss -> local, and says so
        model = tmp_path / "models" / "autoAI" / "xgb-opt" / "2" / "model.pt"
        assert model.exists()
        train_loss = sum(r["logloss"] for r in res) / 3
        assert train_loss < 0.69                                                        # below ln 2: it learned
        # Predict job on the same shards
        import copy
        pj = copy.deepcopy(job)
        pj["metadata"] = {"name": "xgb-predict", "namespace": ns}
        pj.pop("status", None)
        for spec in pj["spec"]["xgbReplicaSpecs"].values():
            ctr = spec["template"]["spec"]["containers"][0]
            ctr["args"] = ["--job_type=Predict" if a.startswith("--job_type") else a for a in ctr["args"]]
        mgr.api.create(pj)
        conds = _wait(lambda: _terminal(mgr, "XGBoostJob", ns, "xgb-predict"), timeout=180)
        plog = let.pod_log(ns, "xgb-predict-master-0")
        assert conds[-1] == "Succeeded", plog
        pr = _json.loads([l for l in plog.splitlines() if l.startswith("{")][-1])
        assert pr["job_type"] == "Predict" and pr["trees"] == 3 and pr["rows"] == 12000
        assert abs(pr["logloss"] - train_loss) < 1e-4 and pr["accuracy"] > 0.55
    finally:
        let.stop()
        mgr.stop()


def test_node_agent_registers_a_node(capsys):
    """The agent shows up as a Node with its GPU capacity, what is still free, and a heartbeat —
    what `kubectl get nodes` shows for a kubelet."""
    from kubedl_b200.apiserver import rest
    from kubedl_b200.cmd import kubedlctl
    mgr = Manager()
    let = Gpulet(mgr.api, num_gpus=4, log_dir=f"/tmp/kubedl-b200/test-pods-{os.getpid()}", node_name="box-7")
    let.start()
    srv = rest.serve(mgr.api, "127.0.0.1", 0)
    try:
        node = _wait(lambda: mgr.api.list("Node"), timeout=10)[0]
        assert node["metadata"]["name"] == "box-7"
        assert node["status"]["capacity"]["nvidia.com/gpu"] == "4" and node["status"]["allocatable"]["nvidia.com/gpu"] == "4"
        assert node["status"]["conditions"][0]["type"] == "Ready" and node["status"]["conditions"][0]["status"] == "True"
        # a pod that asks for a GPU shows up in the node's books at once
        mgr.api.create({"kind": "Pod", "metadata": {"name": "g", "namespace": "d"},
                        "spec": {"restartPolicy": "Never", "containers": [{
                            "name": "c", "image": "x", "command": [sys.executable, "-c", "import time; time.sleep(2)"],
                            "resources": {"limits": {"nvidia.com/gpu": 1}}}]}})
        _wait(lambda: mgr.api.get("Node", "default", "box-7")["status"]["allocatable"]["nvidia.com/gpu"] == "3", timeout=10)
        import json as _json
        assign = _json.loads(mgr.api.get("Node", "default", "box-7")["metadata"]["annotations"]["kubedl-b200/gpu-assignments"])
        assert assign == {"d/g": 0}
        assert kubedlctl.main(["-s", srv.url.split("//")[1], "get", "nodes"]) == 0
        out = capsys.readouterr().out
        assert "box-7" in out and "Ready" in out and "3/4" in out
        _wait(lambda: mgr.api.get("Node", "default", "box-7")["status"]["allocatable"]["nvidia.com/gpu"] == "4", timeout=15)
    finally:
        srv.shutdown()
        let.stop()
        mgr.stop()


def test_two_node_agents_share_one_api_server(tmp_path):
    """--multi-node: two agents bid for every job with a Lease; each job runs exactly once, with all
    its pods on the box that won it."""
    mgr = Manager()
    gang_schedule.register_gang_schedulers(mgr)
    setup_with_manager(mgr, JobControllerConfiguration(), "*", CollectorRegistry())
    a = Gpulet(mgr.api, num_gpus=0, log_dir=str(tmp_path / "a"), node_name="box-a", multi_node=True)
    b = Gpulet(mgr.api, num_gpus=0, log_dir=str(tmp_path / "b"), node_name="box-b", multi_node=True)
    mgr.start()
    a.start()
    b.start()
    try:
        t = lambda: tmpl("tensorflow", image="kubedl-b200/tf-lifecycle-stub",
                         env=[{"name": "KUBEDL_TF_TRAIN_SECONDS", "value": "0.3"}])
        names = [f"mn{i}" for i in range(6)]
        for n in names:
            mgr.api.create(k.new_job("TFJob", n, "mn", {"Worker": {"replicas": 3, "restartPolicy": "Never", "template": t()}},
                                     cleanPodPolicy="None"))
        for n in names:
            conds = _wait(lambda n=n: _terminal(mgr, "TFJob", "mn", n), timeout=120)
            assert conds[-1] == "Succeeded", (n, conds)
        assert a.launched + b.launched == 18                       # every pod started exactly once
        holders = {}
        for n in names:
            nodes = {p["spec"].get("nodeName") for p in mgr.api.list("Pod", "mn", {c.JOB_NAME_LABEL: n})}
            assert len(nodes) == 1 and nodes <= {"box-a", "box-b"}, (n, nodes)     # a job stays on one box
            lease = mgr.api.get("Lease", "mn", f"node-binding-{n}")
            assert lease["spec"]["holderIdentity"] == next(iter(nodes))
            holders[n] = lease["spec"]["holderIdentity"]
        # pods of a box have their logs on that box only
        for n, node in holders.items():
            mine, other = (a, b) if node == "box-a" else (b, a)
            assert "TF_CONFIG ok" in mine.pod_log("mn", f"{n}-worker-0") and other.pod_log("mn", f"{n}-worker-0") == ""
        # the binding goes away with the job
        mgr.api.delete("TFJob", "mn", names[0])
        _wait(lambda: not [l for l in mgr.api.list("Lease", "mn") if c.name_of(l) == f"node-binding-{names[0]}"], timeout=10)
        assert {c.name_of(nd) for nd in mgr.api.list("Node")} == {"box-a", "box-b"}
    finally:
        a.stop()
        b.stop()
        mgr.stop()


def test_pod_thread_budget():
    """One intra-op thread per pod unless the container has a CPU limit (what torchrun does)."""
    from kubedl_b200.node.gpulet import _cpu_threads
    assert _cpu_threads({}) == 1
    assert _cpu_threads({"resources": {"limits": {"cpu": "1500m"}}}) == 2
    assert _cpu_threads({"resources": {"limits": {"cpu": 4}}}) == 4
    assert _cpu_threads({"resources": {"limits": {"cpu": "garbage"}}}) == 1
    mgr = Manager()
    let = Gpulet(mgr.api, num_gpus=0, log_dir=f"/tmp/kubedl-b200/test-pods-{os.getpid()}")
    pod = {"kind": "Pod", "metadata": {"name": "p", "namespace": "d", "labels": {}}, "spec": {}}
    saved = os.environ.pop("OMP_NUM_THREADS", None)
    try:
        assert let.build_env(pod, {"name": "c"}, None)["OMP_NUM_THREADS"] == "1"
        assert let.build_env(pod, {"name": "c", "resources": {"limits": {"cpu": "3"}}}, None)["OMP_NUM_THREADS"] == "3"
        assert let.build_env(pod, {"name": "c", "env": [{"name": "OMP_NUM_THREADS", "value": "7"}]}, None)["OMP_NUM_THREADS"] == "7"
    finally:
        if saved is not None:
            os.environ["OMP_NUM_THREADS"] = saved


def test_node_agent_metrics():
    reg = CollectorRegistry()
    mgr = Manager()
    let = Gpulet(mgr.api, num_gpus=2, log_dir=f"/tmp/kubedl-b200/test-pods-{os.getpid()}", node_name="m0", metrics_registry=reg)
    let.start()
    try:
        assert reg.get_sample_value("kubedl_gpulet_gpus", {"node": "m0", "state": "total"}) == 2
        mgr.api.create({"kind": "Pod", "metadata": {"name": "g", "namespace": "d"},
                        "spec": {"restartPolicy": "Never", "containers": [{
                            "name": "c", "image": "x", "command": [sys.executable, "-c", "import time; time.sleep(1.5)"],
                            "resources": {"limits": {"nvidia.com/gpu": 1}}}]}})
        _wait(lambda: reg.get_sample_value("kubedl_gpulet_pods", {"node": "m0", "state": "running"}) == 1, timeout=10)
        assert reg.get_sample_value("kubedl_gpulet_gpus", {"node": "m0", "state": "free"}) == 1
        assert reg.get_sample_value("kubedl_gpulet_pods_launched_total", {"node": "m0", "how": "exec"}) == 1
        assert reg.get_sample_value("kubedl_gpulet_pod_start_seconds_count", {"node": "m0"}) == 1
        _wait(lambda: reg.get_sample_value("kubedl_gpulet_gpus", {"node": "m0", "state": "free"}) == 2, timeout=15)
        # a second agent on the same registry must not take the process down
        Gpulet(mgr.api, num_gpus=0, log_dir=f"/tmp/kubedl-b200/test-pods-{os.getpid()}", node_name="m1", metrics_registry=reg)
    finally:
        let.stop()
        mgr.stop()


def test_one_status_write_per_pod_start():
    """A pod without init containers goes ADDED -> Running -> Succeeded: no intermediate Pending write
    (each extra pod event is an extra reconcile while the operator is still creating the job's pods)."""
    mgr = Manager()
    let = Gpulet(mgr.api, num_gpus=0, log_dir=f"/tmp/kubedl-b200/test-pods-{os.getpid()}")
    phases = []
    mgr.api.watch("Pod", lambda ev, obj, old: phases.append((ev, (obj.get("status") or {}).get("phase"))))
    let.start()
    try:
        mgr.api.create({"kind": "Pod", "metadata": {"name": "quick", "namespace": "d"},
                        "spec": {"restartPolicy": "Never", "containers": [{
                            "name": "c", "image": "x", "command": [sys.executable, "-c", "pass"]}]}})
        _wait(lambda: phases and phases[-1][1] == "Succeeded", timeout=20)
        assert phases == [("ADDED", None), ("MODIFIED", "Running"), ("MODIFIED", "Succeeded")]
        pod = mgr.api.get("Pod", "d", "quick")
        kinds = {cd["type"] for cd in pod["status"]["conditions"]}
        assert {"PodScheduled", "Initialized", "Ready"} <= kinds and pod["spec"]["nodeName"] == let.node_name
    finally:
        let.stop()
        mgr.stop()


def _pid_alive(pid: int) -> bool:
    try:
        os.kill(pid, 0)
    except ProcessLookupError:
        return False
    try:      # a zombie still answers kill(0)
        with open(f"/proc/{pid}/stat") as f:
            return f.read().split(")")[-1].split()[0] != "Z"
    except FileNotFoundError:
        return False


def test_pod_deleted_during_slow_init_container_leaves_no_orphan(tmp_path):
    """A pod deleted while its launch is in flight (init container still running) must not end up as a
    worker process on a GPU that was already handed to the next pod: the launch is abandoned, whatever it
    started is killed, the GPU returns to the pool and the job's bookkeeping is dropped."""
    mgr = Manager()
    let = Gpulet(mgr.api, num_gpus=1, default_gpu_per_pod=True, log_dir=str(tmp_path / "pods"))
    let.start()
    started, main_ran = tmp_path / "init-started", tmp_path / "main-ran"
    init_cmd = [sys.executable, "-c",
                f"import time,os\nopen(r'{started}','w').write(str(os.getpid()))\ntime.sleep(30)"]
    main_cmd = [sys.executable, "-c", f"open(r'{main_ran}','w').close()\nimport time\ntime.sleep(30)"]
    try:
        mgr.api.create({"kind": "Pod", "metadata": {"name": "slow", "namespace": "d",
                                                    "labels": {c.JOB_NAME_LABEL: "j1"}},
                        "spec": {"restartPolicy": "Never",
                                 "initContainers": [{"name": "init", "image": "x", "command": init_cmd}],
                                 "containers": [{"name": "c", "image": "x", "command": main_cmd}]}})
        _wait(lambda: started.exists() and started.read_text(), timeout=20)
        init_pid = int(started.read_text())
        assert let.free_gpus == [] and ("d", "slow") in let._launching
        mgr.api.delete("Pod", "d", "slow")
        _wait(lambda: ("d", "slow") not in let._launching, timeout=20)
        _wait(lambda: not _pid_alive(init_pid), timeout=10)
        assert not main_ran.exists(), "the worker was started for a pod that no longer exists"
        assert let.free_gpus == [0] and not let._procs and not let._gpu_of
        assert ("d", "slow") not in let._job_of and not let._cancelled
        # the freed GPU serves the next pod, and nothing of the old one is left on it
        mgr.api.create({"kind": "Pod", "metadata": {"name": "next", "namespace": "d"},
                        "spec": {"restartPolicy": "Never", "containers": [{
                            "name": "c", "image": "x", "command": [sys.executable, "-c", "pass"]}]}})
        _wait(lambda: (mgr.api.get("Pod", "d", "next").get("status") or {}).get("phase") == "Succeeded", timeout=20)
        time.sleep(0.2)
        assert not main_ran.exists()
    finally:
        let.stop()
        mgr.stop()


def test_deleted_pod_gpu_returns_only_after_its_process_is_gone(tmp_path):
    """Deleting a running pod signals its process; the GPU goes back to the pool when the process has
    really exited (a worker that takes its grace period still holds device memory)."""
    mgr = Manager()
    let = Gpulet(mgr.api, num_gpus=1, default_gpu_per_pod=True, log_dir=str(tmp_path / "pods"))
    let.start()
    stubborn = [sys.executable, "-c",
                "import signal,time\nsignal.signal(signal.SIGTERM, lambda *a: None)\ntime.sleep(30)"]
    try:
        mgr.api.create({"kind": "Pod", "metadata": {"name": "stubborn", "namespace": "d"},
                        "spec": {"restartPolicy": "Never", "terminationGracePeriodSeconds": 1,
                                 "containers": [{"name": "c", "image": "x", "command": stubborn}]}})
        _wait(lambda: ("d", "stubborn") in let._procs, timeout=20)
        time.sleep(0.3)     # let the interpreter install its SIGTERM handler
        pid = let._procs[("d", "stubborn")].popen.pid
        mgr.api.delete("Pod", "d", "stubborn")
        time.sleep(0.3)
        assert _pid_alive(pid) and let.free_gpus == [], "GPU handed back while the old worker is still alive"
        _wait(lambda: let.free_gpus == [0], timeout=10)      # SIGKILL after the 1 s grace period
        assert not _pid_alive(pid) and not let._terminating
    finally:
        let.stop()
        mgr.stop()


def test_crash_loop_backoff_paces_in_place_restarts(tmp_path):
    """restartPolicy Always/OnFailure: first restart at once, then exponentially spaced, with the
    container reported `waiting: CrashLoopBackOff` in between (the kubelet's behaviour; without it a
    worker dying at import respawned on every 10 ms loop pass)."""
    mgr = Manager()
    let = Gpulet(mgr.api, num_gpus=0, log_dir=str(tmp_path / "pods"))
    let.restart_backoff = (0.4, 5.0, 600.0)
    writes = []
    mgr.api.watch("Pod", lambda ev, obj, old: writes.append(
        ((obj.get("status") or {}).get("containerStatuses") or [{}])[0]))
    let.start()
    try:
        mgr.api.create({"kind": "Pod", "metadata": {"name": "crashy", "namespace": "d"},
                        "spec": {"restartPolicy": "OnFailure", "containers": [{
                            "name": "c", "image": "x", "command": [sys.executable, "-c", "import sys; sys.exit(3)"]}]}})
        t0 = time.time()
        _wait(lambda: any((w.get("state") or {}).get("waiting", {}).get("reason") == "CrashLoopBackOff"
                          for w in writes), timeout=20)
        pod = mgr.api.get("Pod", "d", "crashy")
        assert pod["status"]["phase"] == "Running"
        cs = pod["status"]["containerStatuses"][0]
        assert cs["lastState"]["terminated"]["exitCode"] == 3 and cs["ready"] is False
        time.sleep(2.0)
        restarts = let._procs[("d", "crashy")].restart_count
        # 0 s, +0.4, +0.8, +1.6 ...: a handful of restarts in ~2.5 s, not hundreds
        assert 2 <= restarts <= 5, restarts
        assert len(writes) < 30, len(writes)
        assert time.time() - t0 < 30
    finally:
        let.stop()
        mgr.stop()


def test_gpu_pods_can_be_given_only_their_own_device(monkeypatch):
    """KUBEDL_GPULET_ISOLATE_GPUS=1: a GPU pod's environment makes its GPU the only visible CUDA device (what the
    NVIDIA device plugin does on a real cluster); a mask the user set on the container wins, the agent's own
    mask is honoured; by default every pod sees every device."""
    mgr = Manager()
    ctr = {"name": "c", "image": "x", "command": ["true"]}
    pod = {"kind": "Pod", "metadata": {"name": "p", "namespace": "d", "labels": {}}, "spec": {"containers": [ctr]}}
    monkeypatch.delenv("CUDA_VISIBLE_DEVICES", raising=False)
    monkeypatch.delenv("KUBEDL_GPULET_ISOLATE_GPUS", raising=False)
    let0 = Gpulet(mgr.api, num_gpus=8, log_dir=f"/tmp/kubedl-b200/test-pods-{os.getpid()}")
    env = let0.build_env(pod, ctr, 5)
    assert "CUDA_VISIBLE_DEVICES" not in env and env["KUBEDL_GPU_INDEX"] == "5" and env["LOCAL_RANK"] == "5"
    monkeypatch.setenv("KUBEDL_GPULET_ISOLATE_GPUS", "1")
    let = Gpulet(mgr.api, num_gpus=8, log_dir=f"/tmp/kubedl-b200/test-pods-{os.getpid()}")
    env = let.build_env(pod, ctr, 5)
    assert env["CUDA_VISIBLE_DEVICES"] == "5" and env["KUBEDL_GPU_INDEX"] == "5" and env["LOCAL_RANK"] == "5"
    assert "CUDA_VISIBLE_DEVICES" not in let.build_env(pod, ctr, None)              # a pod without a GPU
    user = dict(ctr, env=[{"name": "CUDA_VISIBLE_DEVICES", "value": "0,1"}])
    assert let.build_env(pod, user, 5)["CUDA_VISIBLE_DEVICES"] == "0,1"
    monkeypatch.setenv("CUDA_VISIBLE_DEVICES", "4,5,6,7")                            # the agent itself is masked
    assert let.build_env(pod, ctr, 2)["CUDA_VISIBLE_DEVICES"] == "6"
