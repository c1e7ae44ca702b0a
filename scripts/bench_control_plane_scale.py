"""Control-plane scaling: cost of one more job (create + reconcile to idle) and of one pod status change with 50 / 200 / 800
live 8-pod jobs in a namespace (fake kubelet, in-process API server).  python scripts/bench_control_plane_scale.py"""
import sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from cp_helpers import Cluster
from kubedl_b200.bench import launch_delay as LD
from kubedl_b200.api import common as c
cl = Cluster()
done = 0
for target in (50, 200, 800):
    t0 = time.perf_counter()
    for i in range(done, target):
        cl.api.create(LD._job("PyTorchJob", f"j{i}", 8, 0.2))
    cl.settle()
    dt = time.perf_counter() - t0
    n = target - done
    print(f"{done:4d} -> {target:4d} jobs live: {dt / n * 1e3:6.2f} ms per job (create + reconcile to idle), pods {len(cl.api.list('Pod', 'bench'))}")
    done = target
t0 = time.perf_counter()
cl.kubelet.set_all("bench", c.POD_RUNNING)
cl.settle()
print("all %d pods -> Running, reconciled: %.2f s (%.2f ms per pod)" % (done * 8, time.perf_counter() - t0, (time.perf_counter() - t0) / (done * 8) * 1e3))
