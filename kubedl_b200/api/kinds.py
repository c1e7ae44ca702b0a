"""The four workload kinds: constants, defaulters and the scheme registry.

One `KindInfo` per CRD carries what the reference spreads over `api/<kind>/<version>/
{constants,defaults,register,types}.go`:

  PyTorchJob  kubeflow.org/v1                    api/pytorch/v1/constants.go:21-36, defaults.go:96-117
  TFJob       kubeflow.org/v1                    api/tensorflow/v1/constants.go:21-34, defaults.go:92-108
  XGBoostJob  xgboostjob.kubeflow.org/v1alpha1   api/xgboost/v1alpha1/constants.go:22-41, defaults.go:37-109
  XDLJob      xdl.kubedl.io/v1alpha1             api/xdl/v1alpha1/constants.go:23-35, defaults.go:37-119

`SCHEME` plays the role of api/apis.go:43-47 + addtoscheme_*.go: every kind registers itself at
import time; `SCHEME.default(obj)` is `scheme.Default(obj)`.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Tuple

from . import common as c


@dataclass
class KindInfo:
    kind: str
    group: str
    version: str
    plural: str
    singular: str
    replica_specs_key: str               # key under .spec holding map[ReplicaType]ReplicaSpec
    replica_types: Tuple[str, ...]
    group_name_label: str                # value of the `group-name` label
    default_container_name: str
    default_port_name: str
    default_port: int
    defaulter: Callable[[dict], None] = field(repr=False, default=lambda job: None)

    @property
    def api_version(self) -> str:
        return f"{self.group}/{self.version}"

    @property
    def crd_name(self) -> str:
        return f"{self.plural}.{self.group}"

    def replica_specs(self, job: dict) -> Dict[str, dict]:
        return job.setdefault("spec", {}).setdefault(self.replica_specs_key, {}) or {}


def set_default_port(pod_spec: dict, container_name: str, port_name: str, port: int) -> None:
    """Add the named default port to the container called `container_name`, else to container 0,
    unless a port of that name is already declared (reference api/pytorch/v1/defaults.go:36-58)."""
    containers = pod_spec.setdefault("containers", [])
    if not containers:
        return
    index = 0
    for i, ctr in enumerate(containers):
        if ctr.get("name") == container_name:
            index = i
            break
    ports = containers[index].get("ports") or []
    if any(p.get("name") == port_name for p in ports):
        return
    ports.append({"name": port_name, "containerPort": port})
    containers[index]["ports"] = ports


def _default_replica(spec: dict, restart_policy: Optional[str]) -> None:
    if spec.get("replicas") is None:
        spec["replicas"] = 1
    if restart_policy is not None and not spec.get("restartPolicy"):
        spec["restartPolicy"] = restart_policy


def _template_spec(spec: dict) -> dict:
    return spec.setdefault("template", {}).setdefault("spec", {})


# ------------------------------------------------------------------ PyTorchJob
PYTORCH_KIND = "PyTorchJob"
PYTORCH_MASTER, PYTORCH_WORKER = "Master", "Worker"
PYTORCH_DEFAULT_PORT_NAME = "pytorchjob-port"
PYTORCH_DEFAULT_CONTAINER = "pytorch"
PYTORCH_DEFAULT_PORT = 23456
PYTORCH_MASTER_RESTART = c.RESTART_POLICY_EXIT_CODE
PYTORCH_WORKER_RESTART = c.RESTART_POLICY_ON_FAILURE


def set_defaults_pytorchjob(job: dict) -> None:
    spec = job.setdefault("spec", {})
    if spec.get("cleanPodPolicy") is None:
        spec["cleanPodPolicy"] = c.CLEAN_POD_POLICY_NONE
    specs = spec.get("pytorchReplicaSpecs") or {}
    c.normalize_keys(specs, (PYTORCH_MASTER, PYTORCH_WORKER))
    for rtype, rs in specs.items():
        if rtype == PYTORCH_WORKER:
            _default_replica(rs, PYTORCH_WORKER_RESTART)
        if rtype == PYTORCH_MASTER:
            _default_replica(rs, PYTORCH_MASTER_RESTART)
            set_default_port(_template_spec(rs), PYTORCH_DEFAULT_CONTAINER, PYTORCH_DEFAULT_PORT_NAME,
                             PYTORCH_DEFAULT_PORT)


# ------------------------------------------------------------------ TFJob
TF_KIND = "TFJob"
TF_PS, TF_WORKER, TF_CHIEF, TF_MASTER, TF_EVAL = "PS", "Worker", "Chief", "Master", "Evaluator"
TF_DEFAULT_PORT_NAME = "tfjob-port"
TF_DEFAULT_CONTAINER = "tensorflow"
TF_DEFAULT_PORT = 2222
TF_DEFAULT_RESTART = c.RESTART_POLICY_EXIT_CODE


def tf_is_chief_or_master(rtype: str) -> bool:
    """reference api/tensorflow/v1/util.go:22-28"""
    return rtype in (TF_CHIEF, TF_MASTER)


def tf_is_worker(rtype: str) -> bool:
    return rtype == TF_WORKER


def tf_is_evaluator(rtype: str) -> bool:
    return rtype == TF_EVAL


def set_defaults_tfjob(job: dict) -> None:
    spec = job.setdefault("spec", {})
    if spec.get("cleanPodPolicy") is None:
        spec["cleanPodPolicy"] = c.CLEAN_POD_POLICY_RUNNING
    specs = spec.get("tfReplicaSpecs") or {}
    c.normalize_keys(specs, (TF_PS, TF_WORKER, TF_CHIEF, TF_MASTER, TF_EVAL))
    for rs in specs.values():
        _default_replica(rs, TF_DEFAULT_RESTART)
        set_default_port(_template_spec(rs), TF_DEFAULT_CONTAINER, TF_DEFAULT_PORT_NAME, TF_DEFAULT_PORT)


# ------------------------------------------------------------------ XGBoostJob
XGB_KIND = "XGBoostJob"
XGB_MASTER, XGB_WORKER = "Master", "Worker"
XGB_DEFAULT_PORT_NAME = "xgboostjob-port"
XGB_DEFAULT_CONTAINER = "xgboostjob"
XGB_DEFAULT_PORT = 9999
XGB_DEFAULT_TTL_SECONDS = 100


def set_defaults_xgboostjob(job: dict) -> None:
    spec = job.setdefault("spec", {})
    if spec.get("cleanPodPolicy") is None:
        spec["cleanPodPolicy"] = c.CLEAN_POD_POLICY_NONE
    if spec.get("ttlSecondsAfterFinished") is None:
        spec["ttlSecondsAfterFinished"] = XGB_DEFAULT_TTL_SECONDS
    specs = spec.get("xgbReplicaSpecs") or {}
    c.normalize_keys(specs, (XGB_MASTER, XGB_WORKER))
    for rs in specs.values():
        _default_replica(rs, None)  # no default restart policy (reference defaults.go:74-78)
        set_default_port(_template_spec(rs), XGB_DEFAULT_CONTAINER, XGB_DEFAULT_PORT_NAME, XGB_DEFAULT_PORT)


# ------------------------------------------------------------------ XDLJob
XDL_KIND = "XDLJob"
XDL_PS, XDL_WORKER, XDL_SCHEDULER, XDL_EXTEND_ROLE = "PS", "Worker", "Scheduler", "ExtendRole"
XDL_DEFAULT_PORT_NAME = "xdljob-port"
XDL_DEFAULT_CONTAINER = "xdl"
XDL_DEFAULT_PORT = 2222
XDL_DEFAULT_RESTART = c.RESTART_POLICY_NEVER
XDL_DEFAULT_MIN_FINISH_RATE = 90
XDL_DEFAULT_BACKOFF_LIMIT = 20


def set_defaults_xdljob(job: dict) -> None:
    spec = job.setdefault("spec", {})
    if spec.get("cleanPodPolicy") is None:
        spec["cleanPodPolicy"] = c.CLEAN_POD_POLICY_RUNNING
    if spec.get("minFinishWorkNum") is None and spec.get("minFinishWorkRate") is None:
        spec["minFinishWorkRate"] = XDL_DEFAULT_MIN_FINISH_RATE
    if spec.get("backoffLimit") is None:
        spec["backoffLimit"] = XDL_DEFAULT_BACKOFF_LIMIT
    specs = spec.get("xdlReplicaSpecs") or {}
    c.normalize_keys(specs, (XDL_PS, XDL_WORKER, XDL_SCHEDULER, XDL_EXTEND_ROLE))
    for rs in specs.values():
        _default_replica(rs, XDL_DEFAULT_RESTART)
        set_default_port(_template_spec(rs), XDL_DEFAULT_CONTAINER, XDL_DEFAULT_PORT_NAME, XDL_DEFAULT_PORT)


# ------------------------------------------------------------------ scheme
class Scheme:
    """kind registry + defaulting (controller-runtime's runtime.Scheme for our purposes)."""

    def __init__(self):
        self._by_kind: Dict[str, KindInfo] = {}
        self._builtin = {"Pod": ("", "v1", "pods"), "Service": ("", "v1", "services"),
                         "Event": ("", "v1", "events"), "ConfigMap": ("", "v1", "configmaps"),
                         "Namespace": ("", "v1", "namespaces"), "Node": ("", "v1", "nodes"),
                         "Deployment": ("apps", "v1", "deployments"),
                         "Lease": ("coordination.k8s.io", "v1", "leases"),
                         "PodGroup": ("scheduling.incubator.k8s.io", "v1alpha1", "podgroups")}

    def register(self, info: KindInfo) -> None:
        self._by_kind[info.kind] = info

    def kinds(self) -> List[str]:
        return list(self._by_kind)

    def workload_kinds(self) -> List[KindInfo]:
        return list(self._by_kind.values())

    def info(self, kind: str) -> KindInfo:
        return self._by_kind[kind]

    def is_workload(self, kind: str) -> bool:
        return kind in self._by_kind

    def knows(self, kind: str) -> bool:
        return kind in self._by_kind or kind in self._builtin

    def default(self, obj: dict) -> dict:
        info = self._by_kind.get(obj.get("kind", ""))
        if info is not None:
            info.defaulter(obj)
        return obj

    def resource(self, kind: str) -> Tuple[str, str, str]:
        """(group, version, plural) of any known kind — what a REST path is built from."""
        if kind in self._by_kind:
            i = self._by_kind[kind]
            return (i.group, i.version, i.plural)
        return self._builtin[kind]

    def all_kinds(self) -> List[str]:
        return list(self._by_kind) + list(self._builtin)

    def api_version(self, kind: str) -> str:
        if kind in self._by_kind:
            return self._by_kind[kind].api_version
        g, v, _ = self._builtin[kind]
        return f"{g}/{v}" if g else v


SCHEME = Scheme()
SCHEME.register(KindInfo(PYTORCH_KIND, "kubeflow.org", "v1", "pytorchjobs", "pytorchjob",
                         "pytorchReplicaSpecs", (PYTORCH_MASTER, PYTORCH_WORKER), "kubeflow.org",
                         PYTORCH_DEFAULT_CONTAINER, PYTORCH_DEFAULT_PORT_NAME, PYTORCH_DEFAULT_PORT,
                         set_defaults_pytorchjob))
SCHEME.register(KindInfo(TF_KIND, "kubeflow.org", "v1", "tfjobs", "tfjob", "tfReplicaSpecs",
                         (TF_PS, TF_WORKER, TF_CHIEF, TF_MASTER, TF_EVAL), "kubeflow.org",
                         TF_DEFAULT_CONTAINER, TF_DEFAULT_PORT_NAME, TF_DEFAULT_PORT, set_defaults_tfjob))
# NB: API group is xgboostjob.kubeflow.org but the label value is kubeflow.org
# (reference api/xgboost/v1alpha1/register.go:33 vs constants.go:24)
SCHEME.register(KindInfo(XGB_KIND, "xgboostjob.kubeflow.org", "v1alpha1", "xgboostjobs", "xgboostjob",
                         "xgbReplicaSpecs", (XGB_MASTER, XGB_WORKER), "kubeflow.org",
                         XGB_DEFAULT_CONTAINER, XGB_DEFAULT_PORT_NAME, XGB_DEFAULT_PORT,
                         set_defaults_xgboostjob))
SCHEME.register(KindInfo(XDL_KIND, "xdl.kubedl.io", "v1alpha1", "xdljobs", "xdljob", "xdlReplicaSpecs",
                         (XDL_PS, XDL_WORKER, XDL_SCHEDULER, XDL_EXTEND_ROLE), "xdl.kubedl.io",
                         XDL_DEFAULT_CONTAINER, XDL_DEFAULT_PORT_NAME, XDL_DEFAULT_PORT, set_defaults_xdljob))


def new_job(kind: str, name: str, namespace: str = "default", replica_specs: Optional[dict] = None,
            **spec_fields) -> dict:
    info = SCHEME.info(kind)
    spec = dict(spec_fields)
    spec[info.replica_specs_key] = replica_specs or {}
    return {"apiVersion": info.api_version, "kind": kind,
            "metadata": {"name": name, "namespace": namespace}, "spec": spec}
