"""`python -m kubedl_b200.cmd.gpulet --master URL` — the node agent as its own process.

What a kubelet is to the reference's operator (it creates Pod objects; something else has to run
them — SURVEY.md §3.6), this is to ours: it watches Pods through the same Kubernetes client the
operator uses, admits gangs all-or-nothing, pins one GPU per pod, starts the containers as local
processes and reports phases, exit codes and restart counts back through the API.
"""
from __future__ import annotations

import argparse
import signal
import sys
import threading

from ..apiserver.kube import KubeAPIServer, KubeConfig
from ..node.gpulet import Gpulet
from ..utils.logger import logger

log = logger("kubedl.gpulet")


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="kubedl-gpulet")
    ap.add_argument("--master", default="", help="URL of the API server")
    ap.add_argument("--kubeconfig", default="")
    ap.add_argument("--context", default=None)
    ap.add_argument("--token-file", default="", help="bearer token for --master")
    ap.add_argument("--gpus", type=int, default=None, help="GPUs to hand out (default: all visible)")
    ap.add_argument("--node-name", default="b200-node-0")
    ap.add_argument("--log-dir", default="/tmp/kubedl-b200/pods")
    ap.add_argument("--prefork-workers", action="store_true", help="fork pods from a launcher that has torch imported")
    ap.add_argument("--warm-gpu-workers", action="store_true",
                    help="keep one pre-forked worker per free GPU with its CUDA context already created")
    ap.add_argument("--multi-node", action="store_true",
                    help="several agents share this API server: bid for each job with a Lease, run only the jobs won")
    args = ap.parse_args(argv)
    from .manager import low_latency_threads
    low_latency_threads()
    api = KubeAPIServer(KubeConfig.load(master=args.master, kubeconfig=args.kubeconfig, context=args.context,
                                        token_file=args.token_file))
    let = Gpulet(api, num_gpus=args.gpus, node_name=args.node_name, log_dir=args.log_dir,
                 prefork=True if args.prefork_workers else None, multi_node=args.multi_node,
                 warm_gpus=True if args.warm_gpu_workers else None)
    let.start()
    print(f"kubedl-b200 gpulet {args.node_name}: {let.num_gpus} GPUs, API {api.config.server}", flush=True)
    stop = threading.Event()
    for sig in (signal.SIGINT, signal.SIGTERM):
        signal.signal(sig, lambda *_: stop.set())
    stop.wait()
    let.stop()
    api.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
