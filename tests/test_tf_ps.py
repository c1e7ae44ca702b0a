"""TFJob worker runtime (`worker/tf_ps_train.py`): parameter-server training driven by nothing
but the TF_CONFIG the operator injects (reference controllers/tensorflow/tensorflow.go:87-141)."""
import json
import os
import socket
import threading
import time

import torch
from prometheus_client import CollectorRegistry

from cp_helpers import load_example, tmpl
from kubedl_b200.api.kinds import new_job
from kubedl_b200.controllers.registry import setup_with_manager
from kubedl_b200.engine.job_controller import JobControllerConfiguration
from kubedl_b200.node.gpulet import Gpulet
from kubedl_b200.plugins import gang_schedule
from kubedl_b200.runtime.manager import Manager
from kubedl_b200.worker import tf_ps_train as ps


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_shard_bounds_cover_everything_once():
    for total, n in ((10, 3), (101_770, 2), (5, 8), (7, 1)):
        b = ps.shard_bounds(total, n)
        assert len(b) == n and b[0][0] == 0 and b[-1][1] == total
        assert all(b[i][1] == b[i + 1][0] for i in range(n - 1))


def test_async_and_sync_application():
    w0 = torch.zeros(4)
    a = ps.ParameterShard(w0, lr=0.5, momentum=0.0, n_workers=2, sync=False)
    a.push(torch.ones(4))
    a.push(torch.ones(4))
    assert a.version == 2 and torch.allclose(a.w, torch.full((4,), -1.0))       # every push applied
    s = ps.ParameterShard(w0, lr=0.5, momentum=0.0, n_workers=2, sync=True)
    s.push(torch.ones(4))
    assert s.version == 0 and torch.equal(s.w, w0)                              # waits for the second worker
    s.push(3 * torch.ones(4))
    assert s.version == 1 and torch.allclose(s.w, torch.full((4,), -1.0))       # mean(1, 3) * lr
    # a finished worker is no longer waited for
    s.push(torch.ones(4))
    assert s.version == 1
    s.worker_done()
    assert s.version == 2
    # momentum: m = mu*m + g
    m = ps.ParameterShard(w0, lr=1.0, momentum=0.5, n_workers=1, sync=False)
    m.push(torch.ones(4))
    m.push(torch.ones(4))
    assert torch.allclose(m.w, torch.full((4,), -2.5))


def test_wire_protocol_pull_push_blocking_pull():
    port = _free_port()
    shard = ps.ParameterShard(torch.arange(6, dtype=torch.float32), lr=1.0, momentum=0.0, n_workers=1, sync=False)
    srv = ps.serve_shard(shard, "127.0.0.1", port)
    try:
        c1, c2 = ps.ShardClient(f"127.0.0.1:{port}"), ps.ShardClient(f"127.0.0.1:{port}")
        v, raw = c1.call(b"PULL", 0)
        assert v == 0 and torch.equal(ps._tensor(raw), torch.arange(6, dtype=torch.float32))
        got = {}
        t = threading.Thread(target=lambda: got.update(r=c2.call(b"PULL", 1)))      # blocks until version 1
        t.start()
        time.sleep(0.1)
        assert not got
        v, _ = c1.call(b"PUSH", 0, torch.ones(6).numpy().tobytes())
        t.join(5)
        assert v == 1 and got["r"][0] == 1
        assert torch.equal(ps._tensor(got["r"][1]), torch.arange(6, dtype=torch.float32) - 1)
        _, stat = c1.call(b"STAT")
        assert json.loads(stat) == {"version": 1, "done": 0, "pushes": 1}
        # the client reconnects when the server side dropped the connection
        c1.sock.close()
        c1.sock = None
        assert c1.call(b"DONE")[0] == 1
        c1.call(b"STOP")
        assert shard.stop.is_set()
    finally:
        srv.shutdown()
        srv.server_close()


def _wait(fn, timeout=180.0):
    t0 = time.time()
    while time.time() - t0 < timeout:
        v = fn()
        if v:
            return v
        time.sleep(0.05)
    raise AssertionError("timed out")


def _cluster(tmp_path):
    mgr = Manager()
    gang_schedule.register_gang_schedulers(mgr)
    recs = setup_with_manager(mgr, JobControllerConfiguration(), "*", CollectorRegistry())
    let = Gpulet(mgr.api, num_gpus=0, log_dir=str(tmp_path / "pods"))
    mgr.start()
    let.start()
    return mgr, let


def _result(log: str) -> dict:
    line = [l for l in log.splitlines() if " result " in l or l.startswith("result ")][-1]
    return json.loads(line.split("result ", 1)[1])


def _run_tfjob(tmp_path, name, sync):
    mgr, let = _cluster(tmp_path)
    try:
        env = [{"name": "KUBEDL_TF_STEPS", "value": "40"}, {"name": "KUBEDL_TF_SYNC", "value": "1" if sync else "0"},
               {"name": "KUBEDL_TF_DEVICE", "value": "cpu"}]
        t = lambda: tmpl("tensorflow", image="kubedl/tf-mnist-with-summaries:1.0", env=list(env),
                         args=["--learning_rate=0.02", "--momentum=0.5", "--batch_size=64", f"--log_dir={tmp_path}/logs"])
        mgr.api.create(new_job("TFJob", name, "kubedl", replica_specs={
            "PS": {"replicas": 2, "template": t()}, "Worker": {"replicas": 2, "template": t()}}))

        def terminal():
            conds = [x["type"] for x in mgr.api.get("TFJob", "kubedl", name).get("status", {}).get("conditions", [])]
            return conds if conds and conds[-1] in ("Succeeded", "Failed") else None
        conds = _wait(terminal)
        logs = {p: let.pod_log("kubedl", f"{name}-{p}") for p in ("ps-0", "ps-1", "worker-0", "worker-1")}
        assert conds == ["Created", "Running", "Succeeded"], logs
        return mgr, logs
    finally:
        let.stop()
        mgr.stop()


def test_tfjob_ps_worker_training_async(tmp_path):
    mgr, logs = _run_tfjob(tmp_path, "psjob", sync=False)
    assert "serving parameters [0, 50885) of 101770" in logs["ps-0"] and "(async, 2 workers)" in logs["ps-0"]
    assert "serving parameters [50885, 101770)" in logs["ps-1"]
    r0 = _result(logs["worker-0"])
    assert r0["steps"] == 40 and r0["last_loss"] < r0["first_loss"]
    assert "eval_loss" in r0                                    # worker 0 is the chief
    # worker 1 may be cleaned up mid-run: the job is decided by worker 0 alone (reference
    # controllers/tensorflow/status.go:86-107) and CleanPodPolicy Running removes what still runs
    if " result " in logs["worker-1"]:
        assert "eval_loss" not in _result(logs["worker-1"])
    assert r0["eval_loss"] < 2.0 and r0["eval_accuracy"] > 0.5
    lines = [json.loads(l) for l in open(tmp_path / "logs" / "summaries.jsonl")]
    assert lines[0]["step"] == 0 and "eval" in lines[-1]
    # the PS replicas never exit by themselves; the operator removed them (CleanPodPolicy Running)
    _wait(lambda: not [p for p in mgr.api.list("Pod", "kubedl") if "-ps-" in p["metadata"]["name"]], timeout=20)


def test_ps_saves_its_shard_when_the_pod_is_deleted(tmp_path):
    """cleanPodPolicy removes the PS pods when the job ends; they get SIGTERM, write a last checkpoint
    and leave, so nothing after the periodic save is lost."""
    mgr, let = _cluster(tmp_path)
    try:
        ck = tmp_path / "ckpt"
        env = [{"name": "KUBEDL_TF_STEPS", "value": "25"}, {"name": "KUBEDL_TF_DEVICE", "value": "cpu"},
               {"name": "KUBEDL_TF_CHECKPOINT_DIR", "value": str(ck)}]
        t = lambda: tmpl("tensorflow", image="kubedl/tf-mnist-with-summaries:1.0", env=list(env),
                         args=["--learning_rate=0.02", "--momentum=0.5", "--checkpoint_every=3600"])
        mgr.api.create(new_job("TFJob", "ck", "kubedl", replica_specs={
            "PS": {"replicas": 1, "template": t()}, "Worker": {"replicas": 1, "template": t()}}))

        def terminal():
            conds = [x["type"] for x in mgr.api.get("TFJob", "kubedl", "ck").get("status", {}).get("conditions", [])]
            return conds if conds and conds[-1] in ("Succeeded", "Failed") else None
        assert _wait(terminal)[-1] == "Succeeded"
        _wait(lambda: (ck / "ps-0.pt").exists(), timeout=20)             # periodic saving was off: this is the SIGTERM save
        state = torch.load(ck / "ps-0.pt", weights_only=False)
        assert state["version"] == 25 and state["pushes"] == 25
        _wait(lambda: "stopping at version 25" in let.pod_log("kubedl", "ck-ps-0"), timeout=10)
    finally:
        let.stop()
        mgr.stop()


def test_tfjob_ps_worker_training_sync(tmp_path):
    _, logs = _run_tfjob(tmp_path, "psync", sync=True)
    assert "(sync, 2 workers)" in logs["ps-0"]
    r0 = _result(logs["worker-0"])
    assert r0["last_loss"] < r0["first_loss"] and r0["eval_loss"] < 2.0


def test_reference_tf_example_runs_single_replica(tmp_path):
    """reference example/tf/tf_job_mnist.yaml: one Worker, no TF_CONFIG (tfjob_controller.go:224-245)."""
    mgr, let = _cluster(tmp_path)
    try:
        (job,) = [d for d in load_example("example/tf/tf_job_mnist.yaml") if d.get("kind") == "TFJob"]
        ctr = job["spec"]["tfReplicaSpecs"]["Worker"]["template"]["spec"]["containers"][0]
        ctr.setdefault("env", []).extend([{"name": "KUBEDL_TF_STEPS", "value": "20"}, {"name": "KUBEDL_TF_DEVICE", "value": "cpu"}])
        mgr.api.create(job)
        ns, name = job["metadata"].get("namespace", "default"), job["metadata"]["name"]

        def terminal():
            conds = [x["type"] for x in mgr.api.get("TFJob", ns, name).get("status", {}).get("conditions", [])]
            return conds if conds and conds[-1] in ("Succeeded", "Failed") else None
        conds = _wait(terminal)
        log = let.pod_log(ns, f"{name}-worker-0")
        assert conds == ["Created", "Running", "Succeeded"], log
        assert "TF_CONFIG absent" in log and _result(log)["steps"] == 20
    finally:
        let.stop()
        mgr.stop()


def test_ps_shard_checkpoint_resume(tmp_path):
    a = ps.ParameterShard(torch.arange(5, dtype=torch.float32), lr=0.1, momentum=0.9, n_workers=1, sync=False)
    a.push(torch.ones(5))
    a.push(torch.ones(5))
    path = str(tmp_path / "ckpt" / "ps-0.pt")
    a.save(path)
    b = ps.ParameterShard(torch.zeros(5), lr=0.1, momentum=0.9, n_workers=1, sync=False)
    assert b.load(path) and b.version == 2 and b.pushes == 2
    a.push(torch.ones(5))
    b.push(torch.ones(5))
    assert torch.equal(a.w, b.w) and torch.equal(a.m, b.m)          # momentum came along
    assert not ps.ParameterShard(torch.zeros(5), 0.1, 0.9, 1, False).load(str(tmp_path / "none.pt"))
    import pytest
    with pytest.raises(ValueError):
        ps.ParameterShard(torch.zeros(7), 0.1, 0.9, 1, False).load(path)
