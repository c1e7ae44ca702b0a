"""Concurrency stress of the control plane (the reference runs `go test` without -race and relies
on one-key-at-a-time work queues) and checkpoint/resume of a restarted pod
(restarted pods keep name/index/rank, which is what lets a worker resume)."""
import os
import threading
import time

from prometheus_client import CollectorRegistry

from cp_helpers import tmpl
from kubedl_b200.api import kinds as k
from kubedl_b200.controllers.registry import setup_with_manager
from kubedl_b200.engine.job_controller import JobControllerConfiguration
from kubedl_b200.node.gpulet import FakeKubelet
from kubedl_b200.runtime.manager import Manager
from kubedl_b200.runtime.workqueue import RateLimitingQueue
from kubedl_b200.utils import status as st


def test_workqueue_never_hands_one_key_to_two_workers():
    q = RateLimitingQueue()
    active, overlaps, done = set(), [], []
    lock = threading.Lock()

    def worker():
        while True:
            item, shut = q.get(block=True, timeout=0.3)
            if item is None:
                return
            with lock:
                if item in active:
                    overlaps.append(item)
                active.add(item)
            time.sleep(0.0005)
            with lock:
                active.discard(item)
                done.append(item)
            q.done(item)

    ts = [threading.Thread(target=worker) for _ in range(8)]
    for t in ts:
        t.start()
    for i in range(3000):
        q.add(i % 17)
    for t in ts:
        t.join()
    assert not overlaps and set(done) == set(range(17))


def test_threaded_manager_many_jobs_converge():
    """40 jobs, 4 reconcile workers per controller, a kubelet thread flipping phases concurrently:
    every job must end Succeeded with exactly its own pods and no duplicates."""
    mgr = Manager()
    recs = setup_with_manager(mgr, JobControllerConfiguration(max_concurrent_reconciles=4), "*", CollectorRegistry())
    kubelet = FakeKubelet(mgr.api)
    mgr.start()
    stop = threading.Event()

    def kubelet_loop():
        while not stop.is_set():
            for pod in mgr.api.list("Pod", "stress"):
                phase = pod.get("status", {}).get("phase")
                try:
                    if not phase:
                        kubelet.set_phase("stress", pod["metadata"]["name"], "Running")
                    elif phase == "Running":
                        kubelet.set_phase("stress", pod["metadata"]["name"], "Succeeded")
                except Exception:
                    pass
            time.sleep(0.005)

    th = threading.Thread(target=kubelet_loop, daemon=True)
    th.start()
    try:
        n = 40
        for i in range(n):
            mgr.api.create(k.new_job("TFJob", f"j{i}", "stress", {
                "Worker": {"replicas": 3, "restartPolicy": "Never", "template": tmpl("tensorflow")}}, cleanPodPolicy="None"))
        deadline = time.time() + 60
        while time.time() < deadline:
            jobs = mgr.api.list("TFJob", "stress")
            if len(jobs) == n and all(st.is_succeeded(j.get("status", {})) for j in jobs):
                break
            time.sleep(0.05)
        jobs = mgr.api.list("TFJob", "stress")
        assert all(st.is_succeeded(j.get("status", {})) for j in jobs), \
            [j["metadata"]["name"] for j in jobs if not st.is_succeeded(j.get("status", {}))]
        pods = mgr.api.list("Pod", "stress")
        assert len(pods) == 3 * n and len({p["metadata"]["name"] for p in pods}) == 3 * n
        assert recs["TFJob"].controller.error_count == 0
    finally:
        stop.set()
        mgr.stop()


def test_worker_checkpoint_resume(tmp_path):
    """mnist worker saves {model, step}; a second run with the same checkpoint continues from it."""
    import subprocess, sys
    ckpt = str(tmp_path / "ckpt.pt")
    env = dict(os.environ, WORLD_SIZE="1", RANK="0", KUBEDL_CHECKPOINT=ckpt, PYTHONPATH=os.getcwd())
    out1 = subprocess.run([sys.executable, "-m", "kubedl_b200.worker.mnist_ddp", "--steps", "10"], env=env,
                          capture_output=True, text=True, timeout=120)
    assert out1.returncode == 0 and "10 steps" in out1.stdout, out1.stdout + out1.stderr
    out2 = subprocess.run([sys.executable, "-m", "kubedl_b200.worker.mnist_ddp", "--steps", "15"], env=env,
                          capture_output=True, text=True, timeout=120)
    assert out2.returncode == 0 and "5 steps" in out2.stdout, out2.stdout + out2.stderr   # resumed at step 10
