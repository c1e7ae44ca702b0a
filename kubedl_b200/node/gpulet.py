"""gpulet — the node agent for one 8×B200 box: scheduler + kubelet + service DNS in one.

The reference stops at "create Pod and Service objects"; a real cluster's scheduler, kubelet and
kube-dns do the rest (SURVEY.md §1 boundaries).  On a single box gpulet plays those parts:

  * scheduling: every pod gets at most one GPU (`nvidia.com/gpu` limit or the job's gang), bound
    by index through KUBEDL_GPU_INDEX / LOCAL_RANK — all 8 devices stay visible by default; with
    KUBEDL_GPULET_ISOLATE_GPUS=1 the pod's GPU is its only visible device (CUDA_VISIBLE_DEVICES): the
    symmetric heap and the NVLS multicast object are shared between pods as POSIX fds of cuMem
    allocations, which need no visible peer devices.  Gang-scheduled pods
    (`schedulerName: gpu-gang` / `kube-batch`) are admitted all-or-nothing per PodGroup.
  * container runtime: an image name maps to a local command (`ImageRegistry`); otherwise the
    container's own command/args run.  `$(VAR)` references are expanded like the kubelet does;
    init containers run first, sequentially (git-sync is emulated with `git clone`).
  * DNS: headless-service names (`<job>-<rtype>-<idx>`, also `<name>.<ns>.svc[...]`) found in env
    values are rewritten to 127.0.0.1; service ports are remapped to free host ports per job so
    several jobs with the same default port can coexist.
  * status: phase, PodScheduled/Initialized/Ready conditions with precise transition times (the
    launch-delay metrics read PodReady), containerStatuses with state/exitCode/restartCount;
    OnFailure/Always restart in place with the kubelet's CrashLoopBackOff (first restart at once,
    then 10 s doubling to 5 min, `state.waiting` in between), Never/ExitCode end the pod.
  * deletion: a pod deleted while its launch is still in flight (init containers) is cancelled by the
    launching thread itself; a running pod's GPU returns to the pool only once its process is gone.
  * fault injection: `kill_pod`, `fail_pod_after` (reference has none).

`FakeKubelet` is the process-free variant for control-plane tests: phases are flipped by hand.
"""
from __future__ import annotations

import json
import os
import re
import shlex
import signal
import socket
import subprocess
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Tuple

from ..api import common as c
from ..apiserver.store import ADDED, DELETED, MODIFIED, APIServer, NotFound
from ..utils.logger import logger

log = logger("kubedl.gpulet")

GANG_SCHEDULERS = ("kube-batch", "gpu-gang")
_VAR = re.compile(r"\$\(([A-Za-z_][A-Za-z0-9_]*)\)")


def expand_vars(s: str, env: Dict[str, str]) -> str:
    """Kubernetes `$(VAR)` expansion; `$$(VAR)` escapes; unknown names are left as is."""
    if "$(" not in s:
        return s
    out, i = [], 0
    while i < len(s):
        if s.startswith("$$(", i):
            j = s.find(")", i)
            if j > 0:
                out.append(s[i + 1:j + 1])
                i = j + 1
                continue
        m = _VAR.match(s, i)
        if m:
            name = m.group(1)
            out.append(env[name] if name in env else m.group(0))
            i = m.end()
        else:
            out.append(s[i])
            i += 1
    return "".join(out)


class _LazyServiceNames:
    """`name in services-of-the-namespace`, listing the services on first use only."""

    def __init__(self, agent, ns: str):
        self._agent, self._ns, self._names = agent, ns, None

    def __contains__(self, name: str) -> bool:
        if self._names is None:
            self._names = self._agent._service_names(self._ns)
        return name in self._names


class ImageRegistry:
    """image name -> local argv.  The stock images of the reference examples are mapped to the
    worker entrypoints of this repo, so the example YAMLs run unchanged."""

    def __init__(self):
        py = sys.executable
        self.images: Dict[str, List[str]] = {
            "kubedl/pytorch-dist-example": [py, "-m", "kubedl_b200.worker.mnist_ddp"],
            "kubedl/tf-mnist-with-summaries:1.0": [py, "-m", "kubedl_b200.worker.tf_ps_train"],
            "kubedl-b200/tf-lifecycle-stub": [py, "-m", "kubedl_b200.worker.tf_stub"],
            "docker.io/merlintang/xgboost-dist-iris:1.1": [py, "-m", "kubedl_b200.worker.xgb_train"],
            "kubedl/xdl-mnist-example": [py, "-m", "kubedl_b200.worker.xdl_train"],
            "kubedl/git-sync:v1": [py, "-m", "kubedl_b200.node.git_sync"],
            "kubedl-b200/resnet50-ddp": [py, "-m", "kubedl_b200.worker.ddp"],
        }

    def register(self, image: str, argv: List[str]) -> None:
        self.images[image] = list(argv)

    def resolve(self, container: dict, env: Dict[str, str]) -> List[str]:
        image = container.get("image", "")
        cmd = [expand_vars(x, env) for x in (container.get("command") or [])]
        args = [expand_vars(x, env) for x in (container.get("args") or [])]
        if image in self.images:
            base = self.images[image]
            # image ENTRYPOINT is replaced by `command`; for mapped images the pod's command was
            # written for the original image filesystem, so the mapped entrypoint wins and the
            # original command line is handed over for argument parsing.
            if cmd:
                return base + ["--original-command", " ".join(shlex.quote(x) for x in cmd)] + args
            return base + args
        if cmd:
            return cmd + args
        raise RuntimeError(f"image {image!r} is not in the local image registry and the container has no command")


def _cpu_threads(container: dict) -> int:
    """Threads for a container: its CPU limit rounded up (`2`, `1500m`), else 1."""
    lim = ((container.get("resources") or {}).get("limits") or {}).get("cpu")
    if lim in (None, ""):
        return 1
    try:
        v = float(str(lim)[:-1]) / 1000.0 if str(lim).endswith("m") else float(lim)
    except ValueError:
        return 1
    return max(1, int(-(-v // 1)))


def free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@dataclass
class _Proc:
    pod_key: Tuple[str, str]
    container: str
    popen: subprocess.Popen
    gpu: Optional[int]
    restart_count: int = 0
    started_at: str = ""
    log_path: str = ""
    uid: str = ""
    # CrashLoopBackOff bookkeeping (restartPolicy Always / OnFailure)
    started_mono: float = 0.0
    crashes: int = 0                 # consecutive crashes, forgotten after a long clean run
    restart_at: Optional[float] = None   # monotonic time of the pending in-place restart
    exit_code: Optional[int] = None


class _LaunchCancelled(Exception):
    """The pod was deleted while its launch was still in flight."""


class Gpulet:
    def __init__(self, api: APIServer, num_gpus: Optional[int] = None, registry: Optional[ImageRegistry] = None,
                 log_dir: str = "/tmp/kubedl-b200/pods", node_name: str = "b200-node-0",
                 default_gpu_per_pod: bool = False, prefork: Optional[bool] = None, multi_node: bool = False,
                 metrics_registry=None, warm_gpus: Optional[bool] = None):
        self.api = api
        self.registry = registry or ImageRegistry()
        if num_gpus is None:
            try:
                import torch
                num_gpus = torch.cuda.device_count()
            except Exception:
                num_gpus = 0
        self.num_gpus = num_gpus
        self.free_gpus: List[int] = list(range(num_gpus))
        self.node_name = node_name
        self.log_dir = log_dir
        self.default_gpu_per_pod = default_gpu_per_pod
        # KUBEDL_GPULET_ISOLATE_GPUS=1: a GPU pod sees only its own device, as under the NVIDIA device plugin on a
        # real cluster.  Validated (all three example jobs at 2 and 8 GPUs); off by default because it buys
        # nothing on this box: time to first step stayed 16 s at 8 pods (profiles/r2_examples_8gpu_isolated.log).
        self.isolate_gpus = os.environ.get("KUBEDL_GPULET_ISOLATE_GPUS", "0") == "1"
        os.makedirs(log_dir, exist_ok=True)
        self._lock = threading.RLock()
        self._procs: Dict[Tuple[str, str], _Proc] = {}
        self._gpu_of: Dict[Tuple[str, str], int] = {}
        self._ports: Dict[Tuple[str, str, int], int] = {}     # (ns, job, declared port) -> host port
        self._pending: Dict[Tuple[str, str], dict] = {}
        # launches in flight: pod key -> pod uid; a DELETED event for such a pod cannot find a process
        # yet, so it leaves a cancellation mark and the launching thread cleans up after itself
        self._launching: Dict[Tuple[str, str], str] = {}
        self._cancelled: set = set()                           # {(key, uid)}
        # deleted pods whose process has been signalled but not yet seen to exit: (proc, gpu)
        self._terminating: List[Tuple["_Proc", Optional[int]]] = []
        # restart back-off for Always / OnFailure, the kubelet's numbers: first restart at once, then
        # 10 s doubling to 5 min, forgotten after 10 min of clean running
        self.restart_backoff = (float(os.environ.get("KUBEDL_GPULET_RESTART_BACKOFF_SECONDS", "10")), 300.0, 600.0)
        self._endpoint_pats: Dict[str, "re.Pattern"] = {}
        self._env_cache = None
        self._stop = threading.Event()
        self._wake = threading.Event()          # set by the watches; the loop otherwise ticks every 10 ms (reaping)
        self._threads: List[threading.Thread] = []
        self._unwatch: List[Callable[[], None]] = []
        self._pool = ThreadPoolExecutor(max_workers=16, thread_name_prefix="gpulet-pod")
        self.launched = 0
        # several node agents on one API server: a job runs where its binding Lease was won
        self.multi_node = multi_node
        self._bindings: Dict[Tuple[str, str], bool] = {}
        self._ready_since = c.now_rfc3339()
        self._node_sig, self._node_at = None, 0.0
        self._start_hist = None
        if metrics_registry is not None:
            try:
                self._register_metrics(metrics_registry)
            except ValueError as e:      # a second agent on the same registry in one process: first one keeps the series
                log.debug("node agent metrics not registered: %s", e)
                self._start_hist = None
        # pre-forked workers (node/zygote.py): pods fork from a process that has imported torch
        if prefork is None:
            prefork = os.environ.get("KUBEDL_GPULET_PREFORK", "0") == "1"
        # warm GPU workers (node/zygote.py): one pre-forked worker per FREE GPU with its CUDA context up and the
        # kernel library loaded, waiting to become the next pod bound to that GPU.  Implies pre-forking.
        if warm_gpus is None:
            warm_gpus = os.environ.get("KUBEDL_GPULET_WARM_GPUS", "0") == "1"
        self.warm_gpus = bool(warm_gpus) and num_gpus > 0
        self.zygote = None
        if prefork or self.warm_gpus:
            from .zygote import ZygoteClient
            self.zygote = ZygoteClient()
            if self.warm_gpus:
                self.zygote.on_ready.append(self._warm_free_gpus)

    # ------------------------------------------------------------------ lifecycle
    def _gpu_env(self, gpu: int) -> Dict[str, str]:
        """What a pod bound to `gpu` gets on top of the agent's environment for device selection."""
        out = {"KUBEDL_GPU_INDEX": str(gpu), "LOCAL_RANK": str(gpu)}
        if self.isolate_gpus:
            # the pod sees ONLY its GPU (workers take device LOCAL_RANK % device_count = 0).  The symmetric
            # heap needs no visible peers: memory is shared as POSIX fds of cuMem allocations and mapped for
            # the local device, the NVLS multicast object likewise (validated with the fused all-reduce check).
            mine = [d for d in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if d.strip()]
            # the agent's own mask, if any, names the devices its indices refer to
            out["CUDA_VISIBLE_DEVICES"] = mine[gpu].strip() if gpu < len(mine) else str(gpu)
        return out

    def _warm_gpu(self, gpu: int) -> None:
        if not self.warm_gpus or self.zygote is None or self._stop.is_set():
            return
        sel = self._gpu_env(gpu)
        cuda_env = {k: v for k, v in sel.items() if k == "CUDA_VISIBLE_DEVICES"}
        self.zygote.warm(gpu, 0 if self.isolate_gpus else gpu, {**os.environ, **sel}, cuda_env)

    def wait_warm(self, timeout: float = 300.0) -> int:
        """Block until a worker is ready on every free GPU (or the timeout); returns how many are."""
        deadline = time.monotonic() + timeout
        while True:
            with self._lock:
                free = list(self.free_gpus)
            ws = self.zygote.warm_workers if self.zygote is not None else {}
            ready = [g for g in free if ws.get(g, {}).get("state") == "ready"]
            if not self.warm_gpus or len(ready) == len(free) or time.monotonic() > deadline:
                return len(ready)
            time.sleep(0.05)

    def _warm_free_gpus(self) -> None:
        with self._lock:
            free = list(self.free_gpus)
        for g in free:
            self._warm_gpu(g)

    def _spawn(self, argv: List[str], env: Dict[str, str], log_path: str, cwd: Optional[str], gpu: Optional[int] = None):
        """Fork from the zygote when it is ready and the command is `python -m module` (or hand the pod to the
        worker already waiting on its GPU), else exec."""
        if self.zygote is not None:
            child = self.zygote.spawn(argv, env, log_path, cwd, gpu=gpu if self.warm_gpus else None)
            if child is not None:
                return child
            if gpu is not None and self.warm_gpus:
                self.zygote.drop(gpu)
        lf = open(log_path, "ab")
        try:
            return subprocess.Popen(argv, env=env, stdout=lf, stderr=subprocess.STDOUT, cwd=cwd, start_new_session=True)
        finally:
            lf.close()

    def start(self) -> None:
        if self.zygote is not None:
            self.zygote.start()
        self._unwatch.append(self.api.watch("Pod", self._on_pod, replay=True, shared=True))     # pods are only read
        self._unwatch.append(self.api.watch("PodGroup", lambda *a: self._kick(), replay=False, shared=True))
        t = threading.Thread(target=self._loop, name="gpulet", daemon=True)
        t.start()
        self._threads.append(t)

    def stop(self) -> None:
        self._stop.set()
        self._wake.set()
        for u in self._unwatch:
            u()
        with self._lock:
            procs = list(self._procs.values()) + [p for p, _ in self._terminating]
        for p in procs:
            self._kill(p)
        self._pool.shutdown(wait=False)
        if self.zygote is not None:
            self.zygote.stop()

    def _kick(self) -> None:
        """Wake the agent's loop now (a pod or PodGroup arrived): admission does not wait for the next poll."""
        self._wake.set()

    def _on_pod(self, ev: str, pod: dict, old: Optional[dict]) -> None:
        key = (c.namespace_of(pod), c.name_of(pod))
        p = None
        with self._lock:
            if ev == DELETED:
                self._pending.pop(key, None)
                uid = self._launching.get(key)
                if uid is not None and uid == c.uid_of(pod):
                    # launch in flight (init containers can take seconds): nothing to signal yet, and the
                    # GPU is not ours to hand back — the launching thread sees the mark, kills whatever
                    # it started and releases the GPU itself
                    self._cancelled.add((key, uid))
                    return
                p = self._procs.get(key)
                if p is not None and p.uid and c.uid_of(pod) and p.uid != c.uid_of(pod):
                    p = None        # a late event for an older incarnation of this name
                else:
                    self._procs.pop(key, None)
                    gpu = self._gpu_of.pop(key, None)
                    if p is not None:
                        # the GPU goes back only when the process is seen gone (_reap): until then the
                        # old worker may still hold its memory
                        self._terminating.append((p, gpu))
                    elif gpu is not None:
                        self._release_gpu(gpu)
                    self._forget_job(key)
            elif ev == ADDED and not (pod.get("status") or {}).get("phase"):
                self._pending[key] = pod
                self._wake.set()
        if ev == DELETED and p is not None:
            self._terminate(p, pod)

    def _release_gpu(self, gpu: Optional[int]) -> None:
        if gpu is not None and gpu not in self.free_gpus:
            self.free_gpus.append(gpu)
            self.free_gpus.sort()
            self._warm_gpu(gpu)              # a worker for the next pod on this GPU starts initialising now

    def _forget_job(self, key: Tuple[str, str]) -> None:
        """Last pod of a job gone: forget the job's host-port assignments.  Caller holds the lock."""
        job = self._job_of.pop(key, None)
        if job and not any(k[0] == key[0] and j == job for k, j in self._job_of.items()):
            for pk in [pk for pk in self._ports
                       if pk[0] == key[0] and (pk[1] == job or str(pk[1]).startswith(job + "-"))]:
                del self._ports[pk]

    # ------------------------------------------------------------------ scheduling
    def _wants_gpu(self, pod: dict) -> bool:
        for ctr in pod.get("spec", {}).get("containers") or []:
            lim = (ctr.get("resources") or {}).get("limits") or {}
            if int(lim.get("nvidia.com/gpu", 0) or 0) > 0:
                return True
        return self.default_gpu_per_pod and self.num_gpus > 0

    def _gang_ready(self, pod: dict, pending: Dict[Tuple[str, str], dict]) -> bool:
        """All-or-nothing: admit a gang-scheduled pod only when minMember pods of its job are
        waiting (or already placed) and enough GPUs are free for all waiting ones."""
        if pod.get("spec", {}).get("schedulerName") not in GANG_SCHEDULERS:
            return True
        job_name = c.labels_of(pod).get(c.JOB_NAME_LABEL, "")
        ns = c.namespace_of(pod)
        try:
            pg = self.api.get("PodGroup", ns, job_name)
        except NotFound:
            return False
        min_member = int(pg.get("spec", {}).get("minMember", 1))
        mates = [p for p in pending.values()
                 if c.namespace_of(p) == ns and c.labels_of(p).get(c.JOB_NAME_LABEL) == job_name]
        with self._lock:
            placed = sum(1 for (pns, pname) in self._procs
                         if pns == ns and self._job_of.get((pns, pname)) == job_name)
        if len(mates) + placed < min_member:
            return False
        need = sum(1 for p in mates if self._wants_gpu(p))
        return need <= len(self.free_gpus)

    @property
    def _job_of(self) -> Dict[Tuple[str, str], str]:
        if not hasattr(self, "_job_of_map"):
            self._job_of_map: Dict[Tuple[str, str], str] = {}
        return self._job_of_map

    def _loop(self) -> None:
        while not self._stop.is_set():
            try:
                self._admit()
                self._reap()
                self._heartbeat()
            except Exception as e:  # keep the node alive
                log.error("gpulet loop error: %s", e)
            self._wake.wait(0.01)
            self._wake.clear()

    # ------------------------------------------------------------------ metrics
    def _register_metrics(self, registry) -> None:
        """What a kubelet's /metrics would tell about this box, next to the operator's kubedl_* series."""
        from prometheus_client import Counter, Gauge, Histogram
        lab = {"node": self.node_name}
        g = Gauge("kubedl_gpulet_gpus", "GPUs managed by the node agent", ["node", "state"], registry=registry)
        g.labels(self.node_name, "total").set_function(lambda: self.num_gpus)
        g.labels(self.node_name, "free").set_function(lambda: len(self.free_gpus))
        pods = Gauge("kubedl_gpulet_pods", "Pods on the node agent", ["node", "state"], registry=registry)
        pods.labels(self.node_name, "running").set_function(lambda: len(self._procs))
        pods.labels(self.node_name, "pending").set_function(lambda: len(self._pending))
        def warm(state):
            z = getattr(self, "zygote", None)
            return sum(1 for w in list(z.warm_workers.values()) if w.get("state") == state) if z is not None else 0
        ww = Gauge("kubedl_gpulet_warm_gpu_workers", "Pre-forked workers holding a CUDA context on a free GPU",
                   ["node", "state"], registry=registry)
        ww.labels(self.node_name, "ready").set_function(lambda: warm("ready"))
        ww.labels(self.node_name, "warming").set_function(lambda: warm("warming"))
        self._launched_total = Counter("kubedl_gpulet_pods_launched_total", "Pods started", ["node", "how"], registry=registry)
        self._start_hist = Histogram("kubedl_gpulet_pod_start_seconds", "Admission to process running, per pod", ["node"],
                                     buckets=(0.002, 0.005, 0.01, 0.025, 0.05, 0.1, 0.25, 0.5, 1, 2.5, 5, 10),
                                     registry=registry).labels(**lab)

    # ------------------------------------------------------------------ Node object
    def node_object(self) -> dict:
        """What a kubelet registers: capacity, what is still allocatable, and a Ready condition whose
        heartbeat shows the agent is alive (`kubedlctl get nodes`)."""
        now = c.now_rfc3339()
        with self._lock:
            free, used = len(self.free_gpus), dict(self._gpu_of)
        return {"apiVersion": "v1", "kind": "Node",
                "metadata": {"name": self.node_name, "namespace": "default",
                             "labels": {"kubernetes.io/hostname": self.node_name, "kubedl-b200/accelerator": "nvidia-b200"},
                             "annotations": {"kubedl-b200/gpu-assignments": json.dumps(
                                 {f"{ns}/{name}": g for (ns, name), g in sorted(used.items())})}},
                "status": {"capacity": {"nvidia.com/gpu": str(self.num_gpus), "cpu": str(os.cpu_count() or 1), "pods": "110"},
                           "allocatable": {"nvidia.com/gpu": str(free), "cpu": str(os.cpu_count() or 1), "pods": "110"},
                           "addresses": [{"type": "InternalIP", "address": "127.0.0.1"},
                                         {"type": "Hostname", "address": self.node_name}],
                           "nodeInfo": {"kubeletVersion": "kubedl-b200-gpulet", "operatingSystem": "linux"},
                           "conditions": [{"type": "Ready", "status": "True", "reason": "GpuletReady",
                                           "message": "node agent is posting ready status",
                                           "lastHeartbeatTime": now, "lastTransitionTime": self._ready_since}]}}

    def _heartbeat(self, force: bool = False) -> None:
        """Register / refresh the Node: every 10 s, and at once when the GPU assignment changed."""
        with self._lock:
            sig = (len(self.free_gpus), tuple(sorted(self._gpu_of.items())))
        now = time.monotonic()
        if not force and sig == self._node_sig and now - self._node_at < 10.0:
            return
        self._node_sig, self._node_at = sig, now
        node = self.node_object()
        try:
            cur = self.api.get("Node", "default", self.node_name)
            node["metadata"]["resourceVersion"] = cur["metadata"].get("resourceVersion")
            self.api.update(node)
        except NotFound:
            try:
                self.api.create(node)
            except Exception as e:
                log.debug("node registration: %s", e)
        except Exception as e:      # a conflict or an API hiccup: the next beat repairs it
            log.debug("node heartbeat: %s", e)
            self._node_at = 0.0

    def _binds_here(self, pod: dict) -> bool:
        """Multi-node: which agent runs a job is decided once per job by creating a Lease
        `node-binding-<job>` (create is atomic: exactly one agent succeeds).  All pods of a job land
        on one box on purpose — the fused collectives need the ranks in one NVSwitch domain, and the
        loopback rendezvous addresses stay valid.  The Lease is owned by the job and goes with it.
        An agent only bids when it could start the pod now (gang complete, GPU free)."""
        if not self.multi_node:
            return True
        ns = c.namespace_of(pod)
        bound = pod.get("spec", {}).get("nodeName")
        if bound:
            return bound == self.node_name
        unit = c.labels_of(pod).get(c.JOB_NAME_LABEL) or f"pod-{c.name_of(pod)}"
        with self._lock:
            if (ns, unit) in self._bindings:
                return self._bindings[(ns, unit)]
        name = f"node-binding-{unit}"
        owner = c.controller_ref(pod)
        lease = {"apiVersion": "coordination.k8s.io/v1", "kind": "Lease",
                 "metadata": {"name": name, "namespace": ns, "labels": {"kubedl-b200/node": self.node_name},
                              "ownerReferences": [owner] if owner else []},
                 "spec": {"holderIdentity": self.node_name, "acquireTime": c.now_precise()}}
        try:
            self.api.create(lease)
            mine = True
        except Exception as e:
            from ..apiserver.store import is_already_exists
            if not is_already_exists(e):
                log.warning("binding %s/%s: %s", ns, unit, e)
                return False            # not cached: try again on the next pass
            try:
                mine = (self.api.get("Lease", ns, name).get("spec") or {}).get("holderIdentity") == self.node_name
            except NotFound:
                return False
        with self._lock:
            self._bindings[(ns, unit)] = mine
            if len(self._bindings) > 20000:
                for k in list(self._bindings)[:10000]:
                    del self._bindings[k]
        return mine

    def _admit(self) -> None:
        with self._lock:
            pending = dict(self._pending)
        for key, pod in sorted(pending.items()):
            if not self._gang_ready(pod, pending):
                continue
            if self.multi_node:
                if self._wants_gpu(pod) and not self.free_gpus:
                    continue          # do not bid for what cannot start here now
                if not self._binds_here(pod):
                    with self._lock:
                        self._pending.pop(key, None)      # another box runs this job
                    continue
            gpu = None
            if self._wants_gpu(pod):
                with self._lock:
                    if not self.free_gpus:
                        continue  # stays Pending until a GPU frees up
                    gpu = self.free_gpus.pop(0)
                    self._gpu_of[key] = gpu
            with self._lock:
                self._pending.pop(key, None)
                self._launching[key] = c.uid_of(pod)
            # pods start concurrently (a kubelet runs one pod worker per pod): the all-pods launch
            # delay is then one process spawn, not eight in a row
            self._pool.submit(self._launch_guarded, key, pod, gpu)

    def _launch_guarded(self, key, pod, gpu) -> None:
        mark = (key, c.uid_of(pod))
        try:
            self._launch(pod, gpu)
        except _LaunchCancelled:
            log.info("pod %s/%s deleted during launch; launch abandoned", *key)
            with self._lock:
                if self._gpu_of.get(key) == gpu:
                    self._gpu_of.pop(key, None)
                self._release_gpu(gpu)
                self._forget_job(key)
        except Exception as e:
            log.warning("launch of %s failed: %s", key, e)
            self._set_status(pod, c.POD_FAILED, reason=str(e), exit_code=127)
            with self._lock:
                if self._gpu_of.get(key) == gpu:
                    self._gpu_of.pop(key, None)
                self._release_gpu(gpu)
        finally:
            with self._lock:
                if self._launching.get(key) == mark[1]:
                    del self._launching[key]
                self._cancelled.discard(mark)

    def _launch_cancelled(self, pod: dict) -> bool:
        with self._lock:
            return ((c.namespace_of(pod), c.name_of(pod)), c.uid_of(pod)) in self._cancelled

    # ------------------------------------------------------------------ env / DNS
    def _service_names(self, ns: str) -> Dict[str, dict]:
        return {c.name_of(s): s for s in self.api.list("Service", ns)}

    def _host_port(self, ns: str, job: str, port: int) -> int:
        with self._lock:
            k = (ns, job, port)
            if k not in self._ports:
                self._ports[k] = free_port()
            return self._ports[k]

    def _rewrite_endpoint(self, value: str, ns: str, job: str, svc_names: Dict[str, dict]) -> str:
        """Replace service DNS names by loopback and declared service ports by host ports."""
        def repl(m):
            host, port = m.group(1), m.group(3)
            short = host.split(".")[0]
            if not host.startswith(job + "-") and short not in svc_names:
                return m.group(0)
            out = "127.0.0.1"
            if port:
                # one host port per (replica, declared port): every entry of a TF_CONFIG cluster map
                # declares the same port on a different DNS name and must stay distinguishable
                out += ":" + str(self._host_port(ns, short, int(port)))
            return out
        if job not in value:
            return value
        pat = self._endpoint_pats.get(job)
        if pat is None:
            if len(self._endpoint_pats) > 256:
                self._endpoint_pats.clear()
            pat = self._endpoint_pats[job] = re.compile(r"\b(" + re.escape(job) + r"-[a-z0-9]+-\d+(\.[A-Za-z0-9.\-]+)?)(?::(\d+))?")
        return pat.sub(repl, value)

    def _base_env(self) -> Dict[str, str]:
        """The agent's own environment as a dict, decoded once and again only when it has changed (building a pod's
        environment used to decode all of os.environ per pod: with eight pods starting at once that and the rest of
        build_env were ~15 ms of interpreter time, most of the all-pods launch delay)."""
        raw = getattr(os.environ, "_data", None)
        if raw is None:
            return dict(os.environ)
        if self._env_cache is None or self._env_cache[0] != raw:
            base = dict(os.environ)
            base["PYTHONPATH"] = os.pathsep.join(filter(None, [os.getcwd(), base.get("PYTHONPATH", "")]))
            self._env_cache = (dict(raw), base)
        return dict(self._env_cache[1])

    def build_env(self, pod: dict, container: dict, gpu: Optional[int]) -> Dict[str, str]:
        ns = c.namespace_of(pod)
        job = c.labels_of(pod).get(c.JOB_NAME_LABEL, "")
        svc_names = _LazyServiceNames(self, ns)      # only a host that does not carry the job's name needs the list
        env = self._base_env()
        declared: Dict[str, str] = {}
        for e in container.get("env") or []:
            if "value" in e:
                v = str(e["value"])
                declared[e["name"]] = expand_vars(v, {**env, **declared}) if "$(" in v else v
        for name, v in list(declared.items()):
            v2 = self._rewrite_endpoint(v, ns, job, svc_names) if job else v
            if name == "MASTER_ADDR" and v2 == "localhost":
                v2 = "127.0.0.1"
            declared[name] = v2
        if "MASTER_PORT" in declared and job:
            declared["MASTER_PORT"] = str(self._host_port(ns, job, int(declared["MASTER_PORT"])))
        env.update(declared)
        if "OMP_NUM_THREADS" not in env:
            # what torchrun does, for the same reason: N pods with one intra-op thread pool per core each
            # oversubscribe the box N-fold.  A CPU limit on the container is honoured instead.
            env["OMP_NUM_THREADS"] = str(_cpu_threads(container))
        env["KUBEDL_POD_NAME"] = c.name_of(pod)
        env["KUBEDL_POD_NAMESPACE"] = ns
        env["KUBEDL_JOB_NAME"] = job
        env["KUBEDL_REPLICA_TYPE"] = c.labels_of(pod).get(c.REPLICA_TYPE_LABEL, "")
        env["KUBEDL_REPLICA_INDEX"] = c.labels_of(pod).get(c.REPLICA_INDEX_LABEL, "")
        if gpu is not None:
            sel = self._gpu_env(gpu)
            if any(e.get("name") == "CUDA_VISIBLE_DEVICES" for e in container.get("env") or []):
                sel.pop("CUDA_VISIBLE_DEVICES", None)        # the pod's own mask wins
            env.update(sel)
        # the worker reports "pod start -> first step" against this (an adopted warm worker's own clock starts
        # at adoption, a fresh process's after the interpreter is up: neither is what the user waits for)
        env["KUBEDL_POD_START_TS"] = repr(time.time())
        return env

    # ------------------------------------------------------------------ launch / status
    def _launch(self, pod: dict, gpu: Optional[int]) -> None:
        t_launch = time.perf_counter()
        key = (c.namespace_of(pod), c.name_of(pod))
        self._job_of[key] = c.labels_of(pod).get(c.JOB_NAME_LABEL, "")
        now = c.now_precise()
        inits = pod.get("spec", {}).get("initContainers") or []
        if inits:
            # only worth a write of its own when something slow follows; otherwise the Running status
            # below carries the same PodScheduled time and the operator sees one pod event, not two
            self._set_status(pod, c.POD_PENDING, conditions=[
                {"type": "PodScheduled", "status": "True", "lastTransitionTime": now}], host=True)
        # init containers, sequentially
        for ic in inits:
            env = self.build_env(pod, ic, gpu)
            argv = self.registry.resolve(ic, env)
            ip = subprocess.Popen(argv, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL,
                                  start_new_session=True)
            while True:
                try:
                    rc = ip.wait(timeout=0.05)
                    break
                except subprocess.TimeoutExpired:
                    if self._launch_cancelled(pod) or self._stop.is_set():
                        try:
                            os.killpg(ip.pid, signal.SIGKILL)      # the exact group started above
                        except (ProcessLookupError, PermissionError):
                            pass
                        ip.wait()
                        raise _LaunchCancelled()
            if rc != 0:
                raise RuntimeError(f"init container {ic.get('name')} exited with {rc}")
        if self._launch_cancelled(pod):
            raise _LaunchCancelled()
        containers = pod.get("spec", {}).get("containers") or []
        if not containers:
            raise RuntimeError("pod has no containers")
        # one process per pod: the default (first) container carries the workload
        ctr = containers[0]
        env = self.build_env(pod, ctr, gpu)
        argv = self.registry.resolve(ctr, env)
        log_path = os.path.join(self.log_dir, f"{key[0]}_{key[1]}.log")
        popen = self._spawn(argv, env, log_path, ctr.get("workingDir") or None, gpu)
        proc = _Proc(key, ctr.get("name", ""), popen, gpu, started_at=c.now_precise(), log_path=log_path,
                     uid=c.uid_of(pod), started_mono=time.monotonic())
        with self._lock:
            # the pod may have been deleted between the check above and the spawn: only a process whose
            # pod is still wanted is registered; the other one is killed here, before its GPU is freed
            cancelled = ((key, proc.uid) in self._cancelled)
            if not cancelled:
                self._procs[key] = proc
        if cancelled:
            self._kill(proc)
            try:
                popen.wait(timeout=10)
            except Exception:
                pass
            raise _LaunchCancelled()
        self.launched += 1
        if self._start_hist is not None:
            self._start_hist.observe(time.perf_counter() - t_launch)
            how = "exec" if isinstance(popen, subprocess.Popen) else ("adopt" if getattr(popen, "adopted", False) else "fork")
            self._launched_total.labels(self.node_name, how).inc()
        ready = c.now_precise()
        self._set_status(pod, c.POD_RUNNING, conditions=[
            {"type": "PodScheduled", "status": "True", "lastTransitionTime": now},
            {"type": "Initialized", "status": "True", "lastTransitionTime": ready},
            {"type": "ContainersReady", "status": "True", "lastTransitionTime": ready},
            {"type": "Ready", "status": "True", "lastTransitionTime": ready}],
            container_statuses=[{"name": proc.container, "ready": True, "restartCount": 0,
                                 "state": {"running": {"startedAt": ready}}}], host=True, start=ready)

    def _set_status(self, pod: dict, phase: str, conditions=None, container_statuses=None, reason: str = "",
                    exit_code: Optional[int] = None, host: bool = False, start: Optional[str] = None) -> None:
        """One merge patch per status change (the kubelet PATCHes pod status too): only the fields named here change,
        whatever else the status holds stays.  It used to be a GET and a PUT — two round trips per pod and transition
        when the API server is remote."""
        status: Dict[str, object] = {"phase": phase}
        if conditions is not None:
            status["conditions"] = conditions
        if container_statuses is not None:
            status["containerStatuses"] = container_statuses
        if reason:
            status["reason"] = reason[:200]
        patch: Dict[str, object] = {"status": status}
        if host:
            status["hostIP"] = "127.0.0.1"
            status["podIP"] = "127.0.0.1"
            patch["spec"] = {"nodeName": self.node_name}
        if start:
            status["startTime"] = start
        if exit_code is not None and container_statuses is None:
            name = (pod.get("spec", {}).get("containers") or [{}])[0].get("name", "")
            status["containerStatuses"] = [{"name": name, "ready": False, "restartCount": 0,
                                            "state": {"terminated": {"exitCode": exit_code, "reason": "Error",
                                                                     "finishedAt": c.now_precise()}}}]
        try:
            self.api.merge_patch("Pod", c.namespace_of(pod), c.name_of(pod), patch)
        except NotFound:
            pass

    def _reap(self) -> None:
        with self._lock:
            procs = list(self._procs.items())
            terminating = list(self._terminating)
        # deleted pods: the GPU goes back to the pool once the signalled process is really gone
        for item in terminating:
            if item[0].popen.poll() is not None:
                with self._lock:
                    if item in self._terminating:
                        self._terminating.remove(item)
                        self._release_gpu(item[1])
        for key, proc in procs:
            if proc.restart_at is not None:         # in CrashLoopBackOff: waiting for its restart time
                if time.monotonic() >= proc.restart_at:
                    try:
                        pod = self.api.get("Pod", key[0], key[1])
                    except NotFound:
                        with self._lock:
                            self._procs.pop(key, None)
                        continue
                    self._restart_in_place(pod, proc)
                continue
            rc = proc.popen.poll()
            if rc is None:
                continue
            if rc < 0:
                rc = 128 - rc  # killed by signal n -> 128+n, like a container runtime reports
            try:
                pod = self.api.get("Pod", key[0], key[1])
            except NotFound:
                with self._lock:
                    self._procs.pop(key, None)
                continue
            policy = pod.get("spec", {}).get("restartPolicy") or "Never"
            if (policy == "Always") or (policy == "OnFailure" and rc != 0):
                self._schedule_restart(pod, proc, rc)
                continue
            with self._lock:
                self._procs.pop(key, None)
                gpu = self._gpu_of.pop(key, None)
                self._release_gpu(gpu)
            done = c.now_precise()
            conds = [cd for cd in (pod.get("status") or {}).get("conditions") or []]
            for cd in conds:
                if cd.get("type") in ("Ready", "ContainersReady"):
                    cd["status"] = "False"
                    cd["lastTransitionTime"] = done
            self._set_status(pod, c.POD_SUCCEEDED if rc == 0 else c.POD_FAILED, conditions=conds,
                             container_statuses=[{"name": proc.container, "ready": False,
                                                  "restartCount": proc.restart_count,
                                                  "state": {"terminated": {"exitCode": rc,
                                                                           "reason": "Completed" if rc == 0 else "Error",
                                                                           "startedAt": proc.started_at,
                                                                           "finishedAt": done}}}])

    def _schedule_restart(self, pod: dict, proc: _Proc, rc: int) -> None:
        """Restart in place (Always / OnFailure) with the kubelet's CrashLoopBackOff: the first restart is
        immediate, later ones wait base * 2^(n-1) seconds up to the cap, and a container that then runs
        cleanly for `reset_after` starts from zero again.  While waiting the pod stays Running with the
        container in `state.waiting` (reason CrashLoopBackOff) — a worker that dies at import no longer
        respawns on every loop pass and floods the API server with status writes."""
        base, cap, reset_after = self.restart_backoff
        now = time.monotonic()
        if proc.started_mono and now - proc.started_mono >= reset_after:
            proc.crashes = 0
        delay = 0.0 if proc.crashes == 0 else min(cap, base * (2.0 ** (proc.crashes - 1)))
        proc.crashes += 1
        proc.exit_code = rc
        if delay <= 0:
            self._restart_in_place(pod, proc)
            return
        proc.restart_at = now + delay
        done = c.now_precise()
        conds = [dict(cd) for cd in (pod.get("status") or {}).get("conditions") or []]
        for cd in conds:
            if cd.get("type") in ("Ready", "ContainersReady"):
                cd["status"] = "False"
                cd["lastTransitionTime"] = done
        self._set_status(pod, c.POD_RUNNING, conditions=conds, container_statuses=[{
            "name": proc.container, "ready": False, "restartCount": proc.restart_count,
            "state": {"waiting": {"reason": "CrashLoopBackOff",
                                  "message": f"back-off {delay:.0f}s restarting failed container"}},
            "lastState": {"terminated": {"exitCode": rc, "reason": "Completed" if rc == 0 else "Error",
                                         "startedAt": proc.started_at, "finishedAt": done}}}])

    def _restart_in_place(self, pod: dict, proc: _Proc) -> None:
        ctr = (pod.get("spec", {}).get("containers") or [{}])[0]
        env = self.build_env(pod, ctr, proc.gpu)
        argv = self.registry.resolve(ctr, env)
        proc.popen = self._spawn(argv, env, proc.log_path, ctr.get("workingDir") or None)
        proc.restart_count += 1
        proc.restart_at = None
        proc.started_at = c.now_precise()
        proc.started_mono = time.monotonic()
        conds = [dict(cd) for cd in (pod.get("status") or {}).get("conditions") or []]
        for cd in conds:
            if cd.get("type") in ("Ready", "ContainersReady") and cd.get("status") != "True":
                cd["status"] = "True"
                cd["lastTransitionTime"] = proc.started_at
        cs = {"name": proc.container, "ready": True, "restartCount": proc.restart_count,
              "state": {"running": {"startedAt": proc.started_at}}}
        if proc.exit_code is not None:
            cs["lastState"] = {"terminated": {"exitCode": proc.exit_code,
                                              "reason": "Completed" if proc.exit_code == 0 else "Error"}}
        self._set_status(pod, c.POD_RUNNING, conditions=conds or None, container_statuses=[cs])

    def _terminate(self, proc: _Proc, pod: dict) -> None:
        """Pod deleted: SIGTERM to its process group, SIGKILL after the grace period if it is still
        there — the kubelet's sequence, so a worker can write its checkpoint on the way out.
        `terminationGracePeriodSeconds` of the pod is honoured (default here 5 s, not 30: a deleted
        pod's GPU is wanted back quickly on a single box)."""
        grace = (pod.get("spec") or {}).get("terminationGracePeriodSeconds")
        grace = float(grace) if grace is not None else float(os.environ.get("KUBEDL_GPULET_GRACE_SECONDS", "5"))
        if grace <= 0:
            return self._kill(proc)
        self._kill(proc, signal.SIGTERM)

        def reap():
            if proc.popen.poll() is None:
                self._kill(proc, signal.SIGKILL)
        t = threading.Timer(grace, reap)
        t.daemon = True
        t.start()

    def _kill(self, proc: _Proc, sig: int = signal.SIGKILL) -> None:
        try:
            os.killpg(proc.popen.pid, sig)   # the exact process group we started
        except (ProcessLookupError, PermissionError):
            pass

    # ------------------------------------------------------------------ fault injection
    def kill_pod(self, namespace: str, name: str, sig: int = signal.SIGKILL) -> bool:
        with self._lock:
            p = self._procs.get((namespace, name))
        if p is None:
            return False
        self._kill(p, sig)
        return True

    def fail_pod_after(self, namespace: str, name: str, seconds: float, sig: int = signal.SIGTERM) -> None:
        threading.Timer(seconds, lambda: self.kill_pod(namespace, name, sig)).start()

    def pod_log(self, namespace: str, name: str) -> str:
        path = os.path.join(self.log_dir, f"{namespace}_{name}.log")
        try:
            with open(path, "r", errors="replace") as f:
                return f.read()
        except FileNotFoundError:
            return ""


class FakeKubelet:
    """Process-free kubelet for control-plane tests: flip pod phases by hand, with the same
    status shape gpulet writes (what reference pkg/test_util/v1/pod.go:57-95 does with
    hand-built Pod objects)."""

    def __init__(self, api: APIServer):
        self.api = api

    def set_phase(self, namespace: str, name: str, phase: str, exit_code: Optional[int] = None,
                  restart_count: int = 0, container: Optional[str] = None, ready_at: Optional[str] = None) -> dict:
        pod = self.api.get("Pod", namespace, name)
        cname = container or (pod["spec"]["containers"][0].get("name", "") if pod["spec"].get("containers") else "")
        status = pod.get("status") or {}
        status["phase"] = phase
        now = ready_at or c.now_precise()
        if phase == c.POD_RUNNING:
            status["conditions"] = [{"type": "PodScheduled", "status": "True", "lastTransitionTime": now},
                                    {"type": "Ready", "status": "True", "lastTransitionTime": now}]
            status["containerStatuses"] = [{"name": cname, "ready": True, "restartCount": restart_count,
                                            "state": {"running": {"startedAt": now}}}]
            status.setdefault("startTime", now)
        elif phase in (c.POD_SUCCEEDED, c.POD_FAILED):
            code = exit_code if exit_code is not None else (0 if phase == c.POD_SUCCEEDED else 1)
            status["containerStatuses"] = [{"name": cname, "ready": False, "restartCount": restart_count,
                                            "state": {"terminated": {"exitCode": code, "finishedAt": now}}}]
        pod["status"] = status
        pod["metadata"].pop("resourceVersion", None)
        return self.api.update(pod)

    def set_all(self, namespace: str, phase: str, selector: Optional[Dict[str, str]] = None, **kw) -> int:
        n = 0
        for pod in self.api.list("Pod", namespace, selector):
            try:
                self.set_phase(namespace, c.name_of(pod), phase, **kw)
            except NotFound:        # deleted since the list (e.g. cleanPodPolicy after the job's last worker finished)
                continue
            n += 1
        return n
