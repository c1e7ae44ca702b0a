"""`python -m kubedl_b200.cmd.manager` — the operator process (reference main.go:48-115).

Order of wiring is the reference's: manager -> schemes -> gang schedulers -> workload controllers
(through the workload gate) -> storage backends -> persist controllers -> metrics HTTP -> start.
Flags are the reference's (main.go:54-60, workload_gate.go:22, persist_controller.go:31-33;
table in reference docs/startup_flags.md) plus what a single-box deployment needs: the REST
façade address, the number of GPUs the node agent manages, and `--no-node-agent`.

Two deployment shapes:
  * all-in-one (default): API server, controllers and node agent in this process;
  * against a cluster: `--master URL`, `--kubeconfig PATH`, `$KUBECONFIG` or the pod's service
    account (reference pkg/util/k8sutil/k8sutil.go:44-69) select a remote Kubernetes API server
    (a real kube-apiserver, or `python -m kubedl_b200.cmd.apiserver`); pods are then run by that
    cluster's kubelets (or by `python -m kubedl_b200.cmd.gpulet`), leader election uses a Lease.
"""
from __future__ import annotations

import argparse
import signal
import sys
import threading
from dataclasses import dataclass, field
from typing import Dict, List, Optional

from ..apiserver import rest
from ..apiserver.kube import KubeAPIServer, KubeConfig
from ..apiserver.store import APIServer
from ..controllers import persist
from ..controllers import registry as controller_registry
from ..engine.job_controller import JobControllerConfiguration
from ..node.gpulet import Gpulet
from ..plugins import controller_metrics, gang_schedule
from ..plugins import metrics as metrics_mod
from ..plugins.storage import backends as storage_backends
from ..runtime.manager import Manager
from ..utils.logger import logger

log = logger("kubedl.setup")

# package-level option holder read by the controllers (reference cmd/options/options.go:22)
CTRL_CONFIG = JobControllerConfiguration()


def _bool(s: str) -> bool:
    return str(s).lower() in ("1", "true", "yes", "y", "")


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(prog="kubedl-manager")
    ap.add_argument("--controller-metrics-addr", default=":8080", help="The address the controller metric endpoint binds to.")
    ap.add_argument("--metrics-addr", type=int, default=8443, help="The address the default endpoints binds to.")
    ap.add_argument("--enable-leader-election", type=_bool, nargs="?", const=True, default=True,
                    help="Enable leader election for controller manager.")
    ap.add_argument("--gang-scheduler-name", default="", help="The scheduler to gang-schedule kubedl jobs, "
                    "gang-scheduling is enabled when it is not empty (kube-batch | gpu-gang).")
    ap.add_argument("--max-reconciles", type=int, default=1, help="The number of max concurrent reconciles of each controller")
    ap.add_argument("--workloads", default="auto", help="List of workloads to be enabled, `*` enables all, `-foo` disables foo.")
    ap.add_argument("--region", default="", help="region tag for persisted rows")
    ap.add_argument("--event-storage", default="", help="event storage backend plugin name, persist events into backend if specified")
    ap.add_argument("--object-storage", default="", help="object storage backend plugin name, persist jobs and pods into backend if specified")
    # single-box additions
    ap.add_argument("--serve", default="127.0.0.1:8001", help="REST façade address for kubedlctl ('' disables)")
    ap.add_argument("--gpus", type=int, default=None, help="GPUs managed by the node agent (default: all visible)")
    ap.add_argument("--no-node-agent", action="store_true", help="control plane only (pods are never started)")
    ap.add_argument("--prefork-workers", action="store_true",
                    help="node agent forks `python -m module` pods from a launcher that has torch imported")
    ap.add_argument("--warm-gpu-workers", action="store_true",
                    help="node agent keeps one pre-forked worker per free GPU with its CUDA context already created "
                         "(implies --prefork-workers)")
    ap.add_argument("--leader-lock", default="/tmp/kubedl-b200-leader.lock")
    # against a cluster
    ap.add_argument("--master", default="", help="URL of a Kubernetes API server to reconcile against "
                    "(overrides the server of the kubeconfig)")
    ap.add_argument("--kubeconfig", default="", help="path to a kubeconfig; $KUBECONFIG and the in-cluster "
                    "service account are tried when --in-cluster is given")
    ap.add_argument("--context", default=None, help="kubeconfig context (default: current-context)")
    ap.add_argument("--token-file", default="", help="bearer token for --master (python -m kubedl_b200.cmd.apiserver --token-file)")
    ap.add_argument("--in-cluster", action="store_true", help="use $KUBECONFIG or the pod's service account")
    ap.add_argument("--leader-election-namespace", default="kubedl-system")
    ap.add_argument("--install-crds", action="store_true", help="create missing CRDs on the remote API server at start")
    ap.add_argument("--pod-create-parallelism", type=int, default=1,
                    help="pods of one replica type created concurrently in slow-start batches (1 = serial, as the reference)")
    return ap


def remote_api_from_args(args) -> Optional[KubeAPIServer]:
    """A `KubeAPIServer` when any of the cluster flags is given, else None (all-in-one)."""
    if not (args.master or args.kubeconfig or args.in_cluster):
        return None
    return KubeAPIServer(KubeConfig.load(master=args.master, kubeconfig=args.kubeconfig, context=args.context,
                                         token_file=getattr(args, "token_file", "")))


@dataclass
class Operator:
    mgr: Manager
    reconcilers: Dict[str, object]
    persisters: Dict[str, object]
    gpulet: Optional[Gpulet]
    servers: List[object] = field(default_factory=list)

    def stop(self) -> None:
        if self.gpulet:
            self.gpulet.stop()
        for s in self.servers:
            try:
                s.shutdown()
            except Exception:
                pass
        self.mgr.stop()


def build_operator(args, api: Optional[APIServer] = None, metrics_registry=None) -> Operator:
    if args.max_reconciles <= 0:
        args.max_reconciles = 1
    CTRL_CONFIG.max_concurrent_reconciles = args.max_reconciles
    CTRL_CONFIG.pod_create_parallelism = max(1, int(getattr(args, "pod_create_parallelism", 1) or 1))
    if args.gang_scheduler_name:
        CTRL_CONFIG.enable_gang_scheduling = True
        CTRL_CONFIG.gang_scheduler_name = args.gang_scheduler_name
    if api is None:
        api = remote_api_from_args(args)
    remote = bool(getattr(api, "remote", False))
    if remote and getattr(args, "install_crds", False):
        from ..api.kinds import SCHEME
        for info in SCHEME.workload_kinds():
            api.install_crd(info.kind)
    mgr = Manager(api=api, leader_election=bool(args.enable_leader_election), leader_lock_path=args.leader_lock,
                  leader_election_namespace=getattr(args, "leader_election_namespace", "kubedl-system"))
    mgr.set_controller_metrics(controller_metrics.ControllerMetrics())
    gang_schedule.register_gang_schedulers(mgr)
    recs = controller_registry.setup_with_manager(mgr, CTRL_CONFIG, args.workloads, metrics_registry)
    storage_backends.register_storage_backends()
    pers = persist.setup_with_manager(mgr, args.object_storage, args.event_storage, args.region, args.workloads)
    # against a cluster the pods belong to its kubelets (or to a separate gpulet process)
    let = None if (args.no_node_agent or remote) else Gpulet(mgr.api, num_gpus=args.gpus,
                                                                  prefork=True if getattr(args, "prefork_workers", False) else None,
                                                                  warm_gpus=True if getattr(args, "warm_gpu_workers", False) else None,
                                                                  metrics_registry=metrics_registry or metrics_mod.DEFAULT_REGISTRY)
    return Operator(mgr, recs, pers, let)


def low_latency_threads() -> None:
    """The operator, the façade and the node agent are threads that hand work to each other many
    times per job (watch event -> queue -> reconcile -> HTTP -> store -> watch ...).  CPython lets
    a running thread keep the GIL for 5 ms before a waiting one gets it; with a few busy threads
    that is 5 ms per hand-off and was most of the all-pods launch delay (measured: 145 -> 84 ms
    over HTTP, `profiles/launch_delay_inproc_vs_http.log`).  None of these threads is throughput
    bound, so trade a little switching overhead for latency."""
    if sys.getswitchinterval() > 0.0005:
        sys.setswitchinterval(0.0005)


def start_operator(op: Operator, args, metrics_registry=None) -> None:
    low_latency_threads()
    if args.metrics_addr:
        try:
            metrics_mod.start_monitoring(int(args.metrics_addr), registry=metrics_registry)
        except OSError as e:
            log.warning("metrics endpoint on %s unavailable: %s", args.metrics_addr, e)
    if op.mgr.controller_metrics is not None:
        try:
            srv = controller_metrics.serve(op.mgr.controller_metrics, args.controller_metrics_addr, manager=op.mgr)
            if srv is not None:
                op.servers.append(srv)
        except OSError as e:
            log.warning("controller metrics endpoint on %s unavailable: %s", args.controller_metrics_addr, e)
    op.mgr.start(block=False)
    if op.gpulet:
        op.gpulet.start()
    if args.serve and not getattr(op.mgr.api, "remote", False):
        host, _, port = args.serve.rpartition(":")
        op.servers.append(rest.serve(op.mgr.api, host or "127.0.0.1", int(port),
                                     log_reader=op.gpulet.pod_log if op.gpulet else None))


def main(argv=None) -> int:
    args = build_parser().parse_args(argv)
    op = build_operator(args)
    start_operator(op, args)
    log.warning("kubedl-b200 manager up: workloads=%s gang=%s rest=%s gpus=%s", list(op.reconcilers),
                args.gang_scheduler_name or "-", args.serve or "-", op.gpulet.num_gpus if op.gpulet else "-")
    stop = threading.Event()
    lost = threading.Event()
    if op.mgr.elector is not None and hasattr(op.mgr.elector, "on_lost"):
        # controller-runtime exits when the lease is lost, and so do we: the Deployment restarts the
        # pod, which queues up behind the new leader instead of lingering as a second writer
        def on_lost():
            lost.set()
            stop.set()
        op.mgr.elector.on_lost = on_lost
    for sig in (signal.SIGINT, signal.SIGTERM):   # reference pkg/util/signals/signal.go
        signal.signal(sig, lambda *_: stop.set())
    stop.wait()
    op.stop()
    if lost.is_set():
        log.error("leader election lost: exiting")
        return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
