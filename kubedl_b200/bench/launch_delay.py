"""Job launch-delay benchmark (second half of BASELINE.json's metric: "job launch delay p50").

Submits N jobs of a kind sequentially to an embedded operator + node agent and reads the raw
observations behind `kubedl_jobs_first_pod_launch_delay_seconds` and
`kubedl_jobs_all_pods_launch_delay_seconds` (definitions: reference pkg/metrics/job_metrics.go:139-194,
docs/metrics.md:16-17), plus time-to-Succeeded.  Pods are real local processes.

    python -m kubedl_b200.bench.launch_delay --jobs 30 --replicas 8
"""
from __future__ import annotations

import argparse
import json
import statistics
import sys
import time

from prometheus_client import CollectorRegistry

from ..api import kinds as k
from ..cmd import manager as M
from ..utils import status as st


def _pct(xs, q):
    xs = sorted(xs)
    if not xs:
        return float("nan")
    i = min(len(xs) - 1, max(0, int(round(q * (len(xs) - 1)))))
    return xs[i]


def _job(kind: str, name: str, replicas: int, seconds: float) -> dict:
    cmd = [sys.executable, "-c", f"import time; time.sleep({seconds})"]
    if kind == "PyTorchJob":
        t = lambda: {"spec": {"containers": [{"name": "pytorch", "image": "python", "command": cmd}]}}
        return k.new_job(kind, name, "bench", {"Master": {"replicas": 1, "restartPolicy": "Never", "template": t()},
                                               "Worker": {"replicas": replicas - 1, "restartPolicy": "Never", "template": t()}},
                         cleanPodPolicy="All")
    if kind == "TFJob":
        t = lambda: {"spec": {"containers": [{"name": "tensorflow", "image": "kubedl/tf-mnist-with-summaries:1.0",
                                              "env": [{"name": "KUBEDL_TF_TRAIN_SECONDS", "value": str(seconds)}]}]}}
        return k.new_job(kind, name, "bench", {"PS": {"replicas": 1, "restartPolicy": "Never", "template": t()},
                                               "Worker": {"replicas": replicas - 1, "restartPolicy": "Never", "template": t()}},
                         cleanPodPolicy="All")
    if kind == "XGBoostJob":
        t = lambda: {"spec": {"containers": [{"name": "xgboostjob", "image": "python", "command": cmd}]}}
        return k.new_job(kind, name, "bench", {"Master": {"replicas": 1, "restartPolicy": "Never", "template": t()},
                                               "Worker": {"replicas": replicas - 1, "restartPolicy": "Never", "template": t()}},
                         cleanPodPolicy="All", ttlSecondsAfterFinished=3600)
    t = lambda: {"spec": {"containers": [{"name": "xdl", "image": "python", "command": cmd}]}}
    return k.new_job(kind, name, "bench", {"Worker": {"replicas": replicas, "restartPolicy": "Never", "template": t()}},
                     cleanPodPolicy="All")


def _spawn_control_plane(prefork: bool):
    """API server and node agent as two processes of their own (`kubedl-b200 apiserver`, `kubedl-b200 gpulet`), the way
    the three components are deployed; returns (url, [stop callables])."""
    import os
    import socket
    import subprocess
    import urllib.request
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    url = f"http://127.0.0.1:{port}"
    env = dict(os.environ, PYTHONPATH=os.pathsep.join(filter(None, [os.getcwd(), os.environ.get("PYTHONPATH", "")])))
    srv = subprocess.Popen([sys.executable, "-m", "kubedl_b200.cmd.apiserver", "--serve", f"127.0.0.1:{port}"],
                           env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    deadline = time.time() + 60
    while True:
        try:
            urllib.request.urlopen(url + "/healthz", timeout=1).read()
            break
        except Exception:
            if time.time() > deadline or srv.poll() is not None:
                srv.kill()
                raise RuntimeError("API server process did not come up")
            time.sleep(0.05)
    let = subprocess.Popen([sys.executable, "-m", "kubedl_b200.cmd.gpulet", "--master", url, "--gpus", "0"]
                           + (["--prefork-workers"] if prefork else []),
                           env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)

    def stop(p):
        def fn():
            p.terminate()                  # the exact process started above
            try:
                p.wait(10)
            except subprocess.TimeoutExpired:
                p.kill()
        return fn
    return url, [stop(let), stop(srv)]


def run(kind: str, jobs: int, replicas: int, seconds: float, max_reconciles: int = 1, remote: bool = False,
        create_parallelism: int = 1, api_latency: float = 0.0, prefork: bool = False, processes: bool = False) -> dict:
    """remote=True puts the API server behind HTTP (`apiserver/rest.py`) with the node agent next to
    it, and makes the operator a client of it (`apiserver/kube.py`) — the reference's deployment
    shape, where every pod / service create and every status write is a round trip."""
    flags = ["--enable-leader-election=false", "--serve", "", "--metrics-addr", "0", "--controller-metrics-addr", "0", "--workloads", "*",
             "--gpus", "0", "--max-reconciles", str(max_reconciles), "--pod-create-parallelism", str(create_parallelism)]
    reg = CollectorRegistry()
    extra = []
    backing = None
    if processes:
        url, extra = _spawn_control_plane(prefork)
        flags += ["--master", url]
        remote = True
    elif remote:
        from ..apiserver import rest
        from ..apiserver.store import APIServer
        from ..node.gpulet import Gpulet
        backing = APIServer()
        srv = rest.serve(backing, "127.0.0.1", 0, latency=api_latency)
        let = Gpulet(backing, num_gpus=0, prefork=prefork)
        let.start()
        extra = [let.stop, srv.shutdown]
        flags += ["--master", srv.url]
    if prefork and not remote:
        flags.append("--prefork-workers")
    args = M.build_parser().parse_args(flags)
    op = M.build_operator(args, metrics_registry=reg)
    M.start_operator(op, args, metrics_registry=reg)
    if prefork and not processes:
        z = (let if remote else op.gpulet).zygote
        z.ready.wait(300)
    if processes:
        time.sleep(8.0 if prefork else 1.0)       # the agent process: watch established (and its launcher up)
    done_times = []
    # the bench's own polling must not load the wire it is measuring: watch the state where it lives
    state = backing if backing is not None else op.mgr.api
    poll = 0.01 if processes else 0.002            # with separate processes every poll is an HTTP GET
    try:
        for i in range(jobs):
            name = f"{kind.lower()}-{i}"
            t0 = time.time()
            op.mgr.api.create(_job(kind, name, replicas, seconds))
            while True:
                s = state.get(kind, "bench", name).get("status", {})
                if st.is_succeeded(s) or st.is_failed(s):
                    break
                if time.time() - t0 > 120:
                    raise RuntimeError(f"{name} did not finish: {s}")
                time.sleep(poll)
            done_times.append(time.time() - t0)
            op.mgr.api.delete(kind, "bench", name)
        m = op.reconcilers[kind].metrics
        first = [v for _, v in m.launch_delays["first_pod"]]
        allp = [v for _, v in m.launch_delays["all_pods"]]
        api = op.mgr.api
        log_ = api.request_log_snapshot() if hasattr(api, "request_log_snapshot") else list(api.request_log)
        creates = sum(1 for verb, kd, _ in log_ if verb == "create" and kd in ("Pod", "Service"))
        return {"kind": kind, "jobs": jobs, "replicas": replicas, "worker_seconds": seconds,
                "first_pod_ms": {"p50": 1e3 * _pct(first, 0.5), "p90": 1e3 * _pct(first, 0.9), "n": len(first)},
                "all_pods_ms": {"p50": 1e3 * _pct(allp, 0.5), "p90": 1e3 * _pct(allp, 0.9), "n": len(allp)},
                "time_to_succeeded_ms": {"p50": 1e3 * _pct(done_times, 0.5), "p90": 1e3 * _pct(done_times, 0.9)},
                "api_creates_per_job": creates / jobs,
                "api": "http, 3 processes" if processes else "http" if remote else "in-process",
                "pod_create_parallelism": create_parallelism, "api_latency_ms": 1e3 * api_latency, "prefork": prefork}
    finally:
        op.stop()
        for fn in extra:
            fn()


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--jobs", type=int, default=30)
    ap.add_argument("--replicas", type=int, default=8)
    ap.add_argument("--seconds", type=float, default=0.2)
    ap.add_argument("--kinds", default="PyTorchJob,TFJob,XGBoostJob,XDLJob")
    ap.add_argument("--remote", action="store_true", help="operator talks to the API server over HTTP")
    ap.add_argument("--processes", action="store_true",
                    help="API server and node agent as separate processes, operator as their HTTP client (implies --remote)")
    ap.add_argument("--max-reconciles", type=int, default=1)
    ap.add_argument("--create-parallelism", type=int, default=1)
    ap.add_argument("--api-latency-ms", type=float, default=0.0, help="with --remote: added to every API request")
    ap.add_argument("--prefork", action="store_true", help="node agent forks `python -m` pods from its launcher")
    a = ap.parse_args(argv)
    for kind in a.kinds.split(","):
        print(json.dumps(run(kind, a.jobs, a.replicas, a.seconds, a.max_reconciles, a.remote, a.create_parallelism,
                             a.api_latency_ms / 1e3, a.prefork, a.processes)), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
