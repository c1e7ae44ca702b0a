"""Node-agent soak: batches of 25 jobs (all four kinds, 3 real processes each) through an embedded operator, printing the
resident set, open descriptors, threads, child processes and the agent's tables after every batch.
    python scripts/soak_node_agent.py"""
import sys, time, os
sys.path.insert(0, ".")
import psutil
from prometheus_client import CollectorRegistry
from kubedl_b200.bench import launch_delay as LD
from kubedl_b200.cmd import manager as M
from kubedl_b200.utils import status as st
flags = ["--enable-leader-election=false", "--serve", "", "--metrics-addr", "0", "--controller-metrics-addr", "0", "--workloads", "*", "--gpus", "0"]
reg = CollectorRegistry()
args = M.build_parser().parse_args(flags)
op = M.build_operator(args, metrics_registry=reg)
M.start_operator(op, args, metrics_registry=reg)
proc = psutil.Process()
let = op.gpulet
def tables():
    names = ["_procs", "_gpu_of", "_ports", "_pending", "_launching", "_cancelled", "_terminating", "_job_of", "_bindings", "_endpoint_pats"]
    return {n: len(getattr(let, n)) for n in names if hasattr(let, n) and len(getattr(let, n))}
try:
    for batch in range(16):
        for i in range(25):
            kind = ("PyTorchJob", "TFJob", "XGBoostJob", "XDLJob")[i % 4]
            name = f"a{batch}-{i}"
            op.mgr.api.create(LD._job(kind, name, 3, 0.02))
            t0 = time.time()
            while True:
                s = op.mgr.api.get(kind, "bench", name).get("status", {})
                if st.is_succeeded(s) or st.is_failed(s): break
                if time.time() - t0 > 60: raise RuntimeError(name)
                time.sleep(0.005)
            op.mgr.api.delete(kind, "bench", name)
        time.sleep(1.0)
        print(batch, "rss %.0f MB" % (proc.memory_info().rss / 1e6), "fds", proc.num_fds(), "threads", proc.num_threads(),
              "children", len(proc.children(recursive=True)), tables(), "pods left", len(op.mgr.api.list("Pod", "bench")), flush=True)
finally:
    op.stop()
