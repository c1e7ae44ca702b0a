"""Rendezvous helpers for the worker entrypoints.

The operator injects framework-specific environments (SURVEY.md §2.5 "Environment contract");
this module turns each of them into (rank, world, master_addr, master_port) for
torch.distributed.  It also repairs the XGBoostJob quirk that master-0 and worker-0 both get
RANK=0 (reference controllers/xgboost/pod.go:106-152, pod_test.go:104-121): the replica type is
read from the pod identity and workers are shifted by one when the job has a master.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Mapping, Optional, Tuple


def replica_identity(env: Mapping[str, str] = os.environ) -> Tuple[str, int]:
    """(replica type lower-case, index) from gpulet's env or from the pod hostname."""
    rt = env.get("KUBEDL_REPLICA_TYPE", "")
    idx = env.get("KUBEDL_REPLICA_INDEX", "")
    if rt and idx != "":
        return rt, int(idx)
    host = env.get("KUBEDL_POD_NAME") or env.get("HOSTNAME", "")
    parts = host.rsplit("-", 2)
    if len(parts) == 3 and parts[2].isdigit():
        return parts[1], int(parts[2])
    return "", int(env.get("RANK", "0"))


def global_rank_from_env(env: Mapping[str, str] = os.environ, kind: str = "pytorch") -> Tuple[int, int]:
    world = int(env.get("WORLD_SIZE", "1"))
    rank = int(env.get("RANK", "0"))
    if kind == "xgboost":
        rt, idx = replica_identity(env)
        if rt == "worker":
            rank = idx + 1  # master-0 is rank 0
        elif rt == "master":
            rank = 0
    return rank, world


def tf_config(env: Mapping[str, str] = os.environ) -> Optional[dict]:
    raw = env.get("TF_CONFIG")
    return json.loads(raw) if raw else None


def tf_rank_world(cfg: dict) -> Tuple[int, int]:
    """Flatten TF_CONFIG's cluster map into a rank (chief/master first, then ps, then workers)."""
    order = ["chief", "master", "ps", "worker"]
    cluster = cfg.get("cluster", {})
    flat = []
    for t in order + sorted(set(cluster) - set(order)):
        flat += [(t, i) for i in range(len(cluster.get(t, [])))]
    me = (cfg["task"]["type"].lower(), int(cfg["task"]["index"]))
    return flat.index(me), len(flat)


def xdl_identity(env: Mapping[str, str] = os.environ) -> Tuple[str, int, str]:
    return env.get("TASK_NAME", ""), int(env.get("TASK_INDEX", "0")), env.get("ZK_ADDR", "")


def init_process_group(backend: str = "gloo", kind: str = "pytorch", env: Mapping[str, str] = os.environ):
    import torch.distributed as dist
    rank, world = global_rank_from_env(env, kind)
    if world > 1 and not dist.is_initialized():
        addr = env.get("MASTER_ADDR", "127.0.0.1")
        if addr == "localhost":
            addr = "127.0.0.1"
        port = env.get("MASTER_PORT", "23456")
        # A rank recreated by the ExitCode policy can reach the previous incarnation's store just
        # before its peers die too; it must give up on that rendezvous (and leave with a retryable
        # code) rather than wait the default half hour.  KUBEDL_RENDEZVOUS_TIMEOUT, seconds.
        import datetime
        timeout = datetime.timedelta(seconds=float(env.get("KUBEDL_RENDEZVOUS_TIMEOUT", "1800")))
        dist.init_process_group(backend, init_method=f"tcp://{addr}:{port}", rank=rank, world_size=world,
                                timeout=timeout)
    return rank, world
