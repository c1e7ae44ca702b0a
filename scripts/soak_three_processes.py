"""Soak of the three-process deployment (API server, node agent, operator over HTTP): batches of 25 jobs with real worker
processes; prints resident set, open descriptors and threads of each process after every batch.  Connections are per thread,
so descriptors / threads of the API server and the agent rise until the agent's 16-thread pool is fully populated, then stay.
    python scripts/soak_three_processes.py"""
import sys, time, os
sys.path.insert(0, ".")
import psutil
from prometheus_client import CollectorRegistry
from kubedl_b200.bench import launch_delay as LD
from kubedl_b200.cmd import manager as M
from kubedl_b200.utils import status as st
url, stops = LD._spawn_control_plane(False)
me = psutil.Process()
kids = {p.pid: p for p in me.children()}
flags = ["--enable-leader-election=false", "--serve", "", "--metrics-addr", "0", "--controller-metrics-addr", "0", "--workloads", "*", "--gpus", "0", "--master", url]
reg = CollectorRegistry()
args = M.build_parser().parse_args(flags)
op = M.build_operator(args, metrics_registry=reg)
M.start_operator(op, args, metrics_registry=reg)
time.sleep(1.5)
def snap():
    out = []
    for p in me.children():
        try:
            out.append("%s rss %.0f MB fds %d thr %d" % (" ".join(p.cmdline()[2:3]).split(".")[-1], p.memory_info().rss / 1e6, p.num_fds(), p.num_threads()))
        except psutil.Error: pass
    return "; ".join(out) + "; operator rss %.0f MB fds %d thr %d" % (me.memory_info().rss / 1e6, me.num_fds(), me.num_threads())
try:
    for batch in range(20):
        for i in range(25):
            kind = ("PyTorchJob", "TFJob", "XGBoostJob", "XDLJob")[i % 4]
            name = f"r{batch}-{i}"
            op.mgr.api.create(LD._job(kind, name, 3, 0.02))
            t0 = time.time()
            while True:
                s = op.mgr.api.get(kind, "bench", name).get("status", {})
                if st.is_succeeded(s) or st.is_failed(s): break
                if time.time() - t0 > 60: raise RuntimeError(name)
                time.sleep(0.01)
            op.mgr.api.delete(kind, "bench", name)
        time.sleep(1.0)
        print(batch, snap(), flush=True)
finally:
    op.stop()
    for s in stops: s()
