"""Control-plane soak: batches of 100 jobs (all four kinds) through create -> Running -> Succeeded -> delete, printing the
resident set and the sizes of the tables that must not grow (expectations, work queues, store, informer cache and index).
    python scripts/soak_control_plane.py"""
import sys, time, os, gc
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import psutil
from cp_helpers import Cluster
from kubedl_b200.bench import launch_delay as LD
from kubedl_b200.api import common as c
cl = Cluster()
proc = psutil.Process()
def sizes():
    out = {}
    for kind, r in cl.recs.items():
        jc = r.ctrl
        exp = getattr(jc, "expectations", None)
        out[kind + ".expectations"] = len(exp) if exp is not None else -1
        bq = getattr(jc, "backoff_queue", None)
        out[kind + ".queue"] = len(r.controller.queue) if hasattr(r.controller.queue, "__len__") else -1
    out["store.objs"] = sum(len(v) for v in cl.api._objs.values())
    out["store.history"] = len(cl.api._history)
    out["cache"] = sum(len(v) for v in cl.mgr.client._cache.values())
    out["index"] = sum(len(v) for v in cl.mgr.client._by_job.values())
    return out
t0 = time.time()
for cycle in range(60):
    for i in range(100):
        name = f"s{cycle}-{i}"
        kind = ("PyTorchJob", "TFJob", "XGBoostJob", "XDLJob")[i % 4]
        cl.api.create(LD._job(kind, name, 4, 0.0))
    cl.settle()
    cl.kubelet.set_all("bench", c.POD_RUNNING); cl.settle()
    cl.kubelet.set_all("bench", c.POD_SUCCEEDED); cl.settle()
    for i in range(100):
        kind = ("PyTorchJob", "TFJob", "XGBoostJob", "XDLJob")[i % 4]
        try: cl.api.delete(kind, "bench", f"s{cycle}-{i}")
        except Exception: pass
    cl.settle()
    cl.api.prune_events(now=time.time() + 7200)
    gc.collect()
    print(cycle, "rss %.0f MB" % (proc.memory_info().rss / 1e6), {k: v for k, v in sizes().items() if v}, "%.0fs" % (time.time() - t0), flush=True)
