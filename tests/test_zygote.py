"""Pre-forked worker launcher (`node/zygote.py`): pods fork from a process that already paid for
the imports.  No counterpart in the reference (kubelet + container runtime start its pods)."""
import os
import signal
import sys
import time

from prometheus_client import CollectorRegistry

from cp_helpers import load_example
from kubedl_b200.controllers.registry import setup_with_manager
from kubedl_b200.engine.job_controller import JobControllerConfiguration
from kubedl_b200.node.gpulet import Gpulet
from kubedl_b200.node.zygote import ZygoteClient
from kubedl_b200.plugins import gang_schedule
from kubedl_b200.runtime.manager import Manager


def _wait(fn, timeout=60.0):
    t0 = time.time()
    while time.time() - t0 < timeout:
        v = fn()
        if v:
            return v
        time.sleep(0.02)
    raise AssertionError("timed out")


def _env(**extra):
    env = dict(os.environ, PYTHONPATH=os.getcwd())
    env.update(extra)
    return env


def test_spawn_exit_codes_logs_and_signals(tmp_path):
    z = ZygoteClient(preload=("json", "kubedl_b200.worker.rendezvous"))
    z.start()
    try:
        assert z.ready.wait(30)
        py = sys.executable
        # not a module run -> the caller has to exec
        assert z.spawn(["/bin/echo", "hi"], _env(), str(tmp_path / "x.log")) is None
        assert not z.eligible([py, "script.py"])
        # a module run: environment and arguments are the pod's, output lands in the pod log
        log = str(tmp_path / "ok.log")
        tfc = '{"cluster": {"worker": ["a:1", "b:2"]}, "task": {"type": "worker", "index": 1}, "environment": "cloud"}'
        p = z.spawn([py, "-m", "kubedl_b200.worker.tf_stub"], _env(TF_CONFIG=tfc, KUBEDL_TF_TRAIN_SECONDS="0.05"), log)
        assert p is not None and p.pid > 0 and p.pid != os.getpid()
        assert p.wait(20) == 0 and p.poll() == 0
        assert "TF_CONFIG ok: task worker:1 rank 1/2" in open(log).read()
        # exit codes pass through
        p = z.spawn([py, "-m", "kubedl_b200.worker.tf_stub"], _env(KUBEDL_FAIL_WITH="3", KUBEDL_TF_TRAIN_SECONDS="0.01"),
                    str(tmp_path / "fail.log"))
        assert p.wait(20) == 3
        # an import error inside the worker is a failure of that pod only
        p = z.spawn([py, "-m", "kubedl_b200.worker.no_such_module"], _env(), str(tmp_path / "bad.log"))
        assert p.wait(20) == 1 and "No module named" in open(tmp_path / "bad.log").read()
        # children lead their own process group: signalling the group reports -signal, like Popen
        p = z.spawn([py, "-m", "kubedl_b200.worker.tf_stub"], _env(KUBEDL_TF_TRAIN_SECONDS="30"), str(tmp_path / "kill.log"))
        _wait(lambda: os.getpgid(p.pid) == p.pid)
        os.killpg(p.pid, signal.SIGTERM)
        assert p.wait(20) == -signal.SIGTERM
        assert z.spawned == 4
        # the environment of one pod does not leak into the next
        p = z.spawn([py, "-m", "kubedl_b200.worker.tf_stub"], _env(KUBEDL_TF_TRAIN_SECONDS="0.01"), str(tmp_path / "clean.log"))
        assert p.wait(20) == 0 and "TF_CONFIG absent" in open(tmp_path / "clean.log").read()
    finally:
        z.stop()
    assert not os.path.exists(z.sock_path)


def test_zygote_death_fails_its_pods(tmp_path):
    z = ZygoteClient(preload=("json",))
    z.start()
    try:
        assert z.ready.wait(30)
        p = z.spawn([sys.executable, "-m", "kubedl_b200.worker.tf_stub"], _env(KUBEDL_TF_TRAIN_SECONDS="30"), str(tmp_path / "a.log"))
        z._proc.kill()                                   # the exact launcher process we started
        assert p.wait(20) == -signal.SIGKILL             # reported as killed; the operator's policy takes over
        assert z.failed.is_set() and not z.ready.is_set()
        assert z.spawn([sys.executable, "-m", "kubedl_b200.worker.tf_stub"], _env(), str(tmp_path / "b.log")) is None
        try:
            os.killpg(p.pid, signal.SIGKILL)             # orphaned by the launcher's death: clean it up
        except ProcessLookupError:
            pass
    finally:
        z.stop()


def test_pytorch_example_through_preforked_workers(tmp_path):
    mgr = Manager()
    gang_schedule.register_gang_schedulers(mgr)
    setup_with_manager(mgr, JobControllerConfiguration(), "*", CollectorRegistry())
    let = Gpulet(mgr.api, num_gpus=0, log_dir=str(tmp_path / "pods"), prefork=True)
    mgr.start()
    let.start()
    try:
        assert let.zygote.ready.wait(120)
        (job,) = load_example("example/pytorch/pytorch_job_mnist_mpi.yaml")
        for spec in job["spec"]["pytorchReplicaSpecs"].values():
            spec["template"]["spec"]["containers"][0]["env"] = [{"name": "KUBEDL_MNIST_STEPS", "value": "5"}]
        t0 = time.time()
        mgr.api.create(job)

        def terminal():
            conds = [x["type"] for x in mgr.api.get("PyTorchJob", "kubedl", "pytorch-dist-sendrecv-example")
                     .get("status", {}).get("conditions", [])]
            return conds if conds and conds[-1] in ("Succeeded", "Failed") else None
        conds = _wait(terminal, timeout=120)
        wall = time.time() - t0
        log = let.pod_log("kubedl", "pytorch-dist-sendrecv-example-worker-1")
        assert conds == ["Created", "Running", "Succeeded"], (conds, log)
        assert "rank 2/3: 5 steps" in log
        assert let.zygote.spawned == 3, let.zygote.spawned           # every pod was a fork of the launcher
        assert wall < 120
    finally:
        let.stop()
        mgr.stop()


def test_warm_worker_is_adopted_not_forked(tmp_path):
    """A worker asked for ahead of time (on a GPU box it would be holding a CUDA context by now) becomes the
    pod bound to its GPU: same pid, the pod's env / argv / log, exit code reported like any other child."""
    z = ZygoteClient(preload=("json", "kubedl_b200.worker.rendezvous"))
    z.start()
    try:
        assert z.ready.wait(30)
        py = sys.executable
        base = _env()
        assert z.warm(0, 0, base)
        w = _wait(lambda: z.warm_workers.get(0) if z.warm_workers.get(0, {}).get("state") == "ready" else None)
        assert w["pid"] > 0 and w["cuda"] is False                       # no GPU here: protocol only
        log = str(tmp_path / "adopted.log")
        p = z.spawn([py, "-m", "kubedl_b200.worker.tf_stub"], _env(KUBEDL_TF_TRAIN_SECONDS="0.05", MARK="from-pod"), log, gpu=0)
        assert p is not None and p.pid == w["pid"] and z.adopted == 1 and 0 not in z.warm_workers
        assert p.wait(20) == 0 and "TF_CONFIG absent" in open(log).read()
        # a request for a GPU with nobody waiting is a plain fork
        p2 = z.spawn([py, "-m", "kubedl_b200.worker.tf_stub"], _env(KUBEDL_FAIL_WITH="4", KUBEDL_TF_TRAIN_SECONDS="0.01"),
                     str(tmp_path / "plain.log"), gpu=0)
        assert p2.wait(20) == 4 and z.adopted == 1
        # adoption of a worker that is still initialising: the request waits in its socket
        assert z.warm(1, 1, base)
        p3 = z.spawn([py, "-m", "kubedl_b200.worker.tf_stub"], _env(KUBEDL_TF_TRAIN_SECONDS="0.01"), str(tmp_path / "early.log"), gpu=1)
        assert p3 is not None and p3.wait(20) == 0 and z.adopted == 2
        # a pod whose CUDA-relevant environment differs from the waiting worker's cannot take it over
        assert z.warm(2, 2, base)
        pid2 = _wait(lambda: z.warm_workers.get(2, {}).get("pid"))
        p4 = z.spawn([py, "-m", "kubedl_b200.worker.tf_stub"], _env(KUBEDL_TF_TRAIN_SECONDS="0.01", CUDA_MODULE_LOADING="EAGER"),
                     str(tmp_path / "other.log"), gpu=2)
        assert p4.pid != pid2 and p4.wait(20) == 0 and z.adopted == 2 and 2 not in z.warm_workers

        def gone(pid):
            try:
                os.kill(pid, 0)
                return False
            except ProcessLookupError:
                return True
        _wait(lambda: gone(pid2))                                        # the dropped worker left
        # a waiting worker that dies is reported and forgotten
        assert z.warm(3, 3, base)
        pid3 = _wait(lambda: z.warm_workers.get(3, {}).get("pid"))
        os.killpg(pid3, signal.SIGKILL)                                  # the exact group forked for GPU 3
        _wait(lambda: 3 not in z.warm_workers)
    finally:
        z.stop()


def test_node_agent_keeps_a_warm_worker_per_free_gpu(tmp_path):
    mgr = Manager()
    gang_schedule.register_gang_schedulers(mgr)
    setup_with_manager(mgr, JobControllerConfiguration(), "*", CollectorRegistry())
    reg = CollectorRegistry()
    let = Gpulet(mgr.api, num_gpus=2, log_dir=str(tmp_path / "pods"), warm_gpus=True, default_gpu_per_pod=True,
                 metrics_registry=reg)
    let.zygote.preload = ("json", "kubedl_b200.worker.rendezvous")
    mgr.start()
    let.start()
    try:
        assert let.zygote.ready.wait(60)
        _wait(lambda: sorted(g for g, w in let.zygote.warm_workers.items() if w["state"] == "ready") == [0, 1])
        first = {g: w["pid"] for g, w in let.zygote.warm_workers.items()}
        assert reg.get_sample_value("kubedl_gpulet_warm_gpu_workers", {"node": let.node_name, "state": "ready"}) == 2
        from kubedl_b200.bench.launch_delay import _job
        mgr.api.create(_job("TFJob", "warm", 2, 0.1))      # PS + worker, both `python -m` workers

        def done():
            s = mgr.api.get("TFJob", "bench", "warm").get("status", {})
            return [x["type"] for x in s.get("conditions", [])][-1:] == ["Succeeded"]
        _wait(done, timeout=60)
        assert let.zygote.adopted == 2 and let.zygote.spawned == 2          # both pods were adoptions
        assert reg.get_sample_value("kubedl_gpulet_pods_launched_total", {"node": let.node_name, "how": "adopt"}) == 2
        # the GPUs came back, and each got a NEW waiting worker
        _wait(lambda: sorted(g for g, w in let.zygote.warm_workers.items() if w["state"] == "ready") == [0, 1])
        again = {g: w["pid"] for g, w in let.zygote.warm_workers.items()}
        assert set(again.values()).isdisjoint(first.values())
    finally:
        let.stop()
        mgr.stop()


def test_forked_pod_honours_its_thread_count(tmp_path):
    """The launcher loaded OpenMP without the pod's OMP_NUM_THREADS; the fork applies it."""
    z = ZygoteClient(preload=("torch",))
    z.start()
    try:
        assert z.ready.wait(120)
        (tmp_path / "show_threads.py").write_text("import torch\nprint('threads', torch.get_num_threads())\n")
        log = str(tmp_path / "t.log")
        p = z.spawn([sys.executable, "-m", "show_threads"], _env(OMP_NUM_THREADS="3", PYTHONPATH=f"{tmp_path}:{os.getcwd()}"), log)
        assert p.wait(60) == 0 and "threads 3" in open(log).read(), open(log).read()
    finally:
        z.stop()


def test_launcher_does_not_outlive_its_node_agent(tmp_path):
    """The agent's connection closing without a goodbye (agent killed) takes the launcher, its pods and its waiting
    workers down after a short grace: nothing is left holding GPU memory with nobody to report to."""
    z = ZygoteClient(preload=("json",))
    z.start()
    try:
        assert z.ready.wait(30)
        p = z.spawn([sys.executable, "-m", "kubedl_b200.worker.tf_stub"], _env(KUBEDL_TF_TRAIN_SECONDS="60"), str(tmp_path / "a.log"))
        assert z.warm(0, 0, _env())
        wpid = _wait(lambda: z.warm_workers.get(0, {}).get("pid"))
        z._sock.shutdown(2)                       # what the launcher sees when the agent process dies
        z._sock.close()
        assert z._proc.wait(15) == 0              # the launcher left by itself

        def gone(pid):
            try:
                os.kill(pid, 0)
                return False
            except ProcessLookupError:
                return True
        _wait(lambda: gone(p.pid) and gone(wpid), timeout=15)
    finally:
        z.stop()
