"""Common job API (v1): the schema shared by all four kinds.

Objects are plain nested dicts in Kubernetes wire format (exactly what the example YAMLs parse
to), so reference manifests load unchanged.  This module is the typed vocabulary around those
dicts: constants, accessors and constructors for ReplicaSpec / RunPolicy / JobStatus /
JobCondition.

Parity: reference pkg/job_controller/api/v1/types.go:23-191 (types),
pkg/job_controller/api/v1/constants.go:3-31 (labels, annotations, namespace).
"""
from __future__ import annotations

import copy
import datetime as _dt
from typing import Any, Dict, Iterable, List, Optional

# ---- label / annotation keys (reference pkg/job_controller/api/v1/constants.go:3-31)
REPLICA_INDEX_LABEL = "replica-index"
REPLICA_TYPE_LABEL = "replica-type"
GROUP_NAME_LABEL = "group-name"
JOB_NAME_LABEL = "job-name"
JOB_ROLE_LABEL = "job-role"
ANNOTATION_GIT_SYNC_CONFIG = "kubedl.io/git-sync-config"
ANNOTATION_TENANCY_INFO = "kubedl.io/tenancy"
KUBEDL_PREFIX = "kubedl.io"
DEFAULT_KUBEDL_NAMESPACE = "kubedl"

# ---- condition types (types.go:103-127)
JOB_CREATED = "Created"
JOB_RUNNING = "Running"
JOB_RESTARTING = "Restarting"
JOB_SUCCEEDED = "Succeeded"
JOB_FAILED = "Failed"
CONDITION_TYPES = (JOB_CREATED, JOB_RUNNING, JOB_RESTARTING, JOB_SUCCEEDED, JOB_FAILED)

CONDITION_TRUE = "True"
CONDITION_FALSE = "False"
CONDITION_UNKNOWN = "Unknown"

# ---- clean pod policy (types.go:129-141)
CLEAN_POD_POLICY_ALL = "All"
CLEAN_POD_POLICY_RUNNING = "Running"
CLEAN_POD_POLICY_NONE = "None"
CLEAN_POD_POLICIES = (CLEAN_POD_POLICY_ALL, CLEAN_POD_POLICY_RUNNING, CLEAN_POD_POLICY_NONE)

# ---- restart policy (types.go:143-156)
RESTART_POLICY_ALWAYS = "Always"
RESTART_POLICY_ON_FAILURE = "OnFailure"
RESTART_POLICY_NEVER = "Never"
RESTART_POLICY_EXIT_CODE = "ExitCode"
RESTART_POLICIES = (RESTART_POLICY_ALWAYS, RESTART_POLICY_ON_FAILURE, RESTART_POLICY_NEVER,
                    RESTART_POLICY_EXIT_CODE)

# ---- pod phases (core/v1)
POD_PENDING = "Pending"
POD_RUNNING = "Running"
POD_SUCCEEDED = "Succeeded"
POD_FAILED = "Failed"
POD_UNKNOWN = "Unknown"


def now_rfc3339(t: Optional[float] = None) -> str:
    d = _dt.datetime.fromtimestamp(t, _dt.timezone.utc) if t is not None else _dt.datetime.now(_dt.timezone.utc)
    return d.strftime("%Y-%m-%dT%H:%M:%S.%fZ") if False else d.strftime("%Y-%m-%dT%H:%M:%SZ")


def now_precise() -> str:
    """RFC3339 with microseconds (used where sub-second launch delays are measured)."""
    return _dt.datetime.now(_dt.timezone.utc).strftime("%Y-%m-%dT%H:%M:%S.%fZ")


def parse_time(s: Optional[str]) -> Optional[float]:
    if not s:
        return None
    try:        # the C parser (3.11+ reads the trailing Z and fractions): ~20x cheaper than strptime, and this runs
        d = _dt.datetime.fromisoformat(s)       # per pod per reconcile and per stored Event when Events are pruned
        if d.tzinfo is None:
            d = d.replace(tzinfo=_dt.timezone.utc)
        return d.timestamp()
    except ValueError:
        pass
    for fmt in ("%Y-%m-%dT%H:%M:%S.%fZ", "%Y-%m-%dT%H:%M:%SZ"):
        try:
            return _dt.datetime.strptime(s, fmt).replace(tzinfo=_dt.timezone.utc).timestamp()
        except ValueError:
            continue
    raise ValueError(f"bad timestamp {s!r}")


# ------------------------------------------------------------------ metadata helpers
def meta(obj: dict) -> dict:
    return obj.setdefault("metadata", {})


def name_of(obj: dict) -> str:
    return obj.get("metadata", {}).get("name", "")


def namespace_of(obj: dict) -> str:
    return obj.get("metadata", {}).get("namespace", "") or "default"


def uid_of(obj: dict) -> str:
    return obj.get("metadata", {}).get("uid", "")


def labels_of(obj: dict) -> Dict[str, str]:
    return obj.get("metadata", {}).get("labels") or {}


def annotations_of(obj: dict) -> Dict[str, str]:
    return obj.get("metadata", {}).get("annotations") or {}


def key_of(obj: dict) -> str:
    """namespace/name, the work-queue key (k8s controller.KeyFunc)."""
    return f"{namespace_of(obj)}/{name_of(obj)}"


def controller_ref(obj: dict) -> Optional[dict]:
    for ref in obj.get("metadata", {}).get("ownerReferences") or []:
        if ref.get("controller"):
            return ref
    return None


# ------------------------------------------------------------------ ReplicaSpec (types.go:65-79)
def new_replica_spec(replicas: Optional[int] = None, template: Optional[dict] = None,
                     restart_policy: Optional[str] = None) -> dict:
    spec: Dict[str, Any] = {"template": template if template is not None else {"spec": {"containers": []}}}
    if replicas is not None:
        spec["replicas"] = replicas
    if restart_policy is not None:
        spec["restartPolicy"] = restart_policy
    return spec


def replicas_of(spec: dict) -> int:
    r = spec.get("replicas")
    return 1 if r is None else int(r)


def total_replicas(specs: Dict[str, dict]) -> int:
    """Σ replicas over all types (reference pkg/util/k8sutil/k8sutil.go:113-120)."""
    return sum(replicas_of(s) for s in (specs or {}).values())


def containers_of(pod_or_template: dict) -> List[dict]:
    return pod_or_template.setdefault("spec", {}).setdefault("containers", [])


# ------------------------------------------------------------------ RunPolicy (types.go:162-191)
RUN_POLICY_FIELDS = ("cleanPodPolicy", "ttlSecondsAfterFinished", "activeDeadlineSeconds",
                     "backoffLimit", "schedulingPolicy")


def run_policy_of(job: dict) -> dict:
    """RunPolicy is inlined in .spec for every kind (json:",inline")."""
    spec = job.get("spec", {})
    return {k: spec[k] for k in RUN_POLICY_FIELDS if k in spec and spec[k] is not None}


# ------------------------------------------------------------------ JobStatus (types.go:23-61)
def new_job_status() -> dict:
    return {"conditions": [], "replicaStatuses": {}}


def status_of(job: dict) -> dict:
    st = job.setdefault("status", {})
    st.setdefault("conditions", [])
    st.setdefault("replicaStatuses", {})
    if st["conditions"] is None:
        st["conditions"] = []
    if st["replicaStatuses"] is None:
        st["replicaStatuses"] = {}
    return st


def new_replica_status() -> dict:
    return {"active": 0, "succeeded": 0, "failed": 0}


def replica_status(status: dict, rtype: str) -> dict:
    rs = status.setdefault("replicaStatuses", {}).get(rtype)
    if rs is None:
        rs = new_replica_status()
        status["replicaStatuses"][rtype] = rs
    for k in ("active", "succeeded", "failed"):
        rs.setdefault(k, 0)
    return rs


def total_active(status: dict) -> int:
    return sum(int(rs.get("active", 0)) for rs in (status.get("replicaStatuses") or {}).values())


def total_failed(status: dict) -> int:
    return sum(int(rs.get("failed", 0)) for rs in (status.get("replicaStatuses") or {}).values())


_SCALARS = (str, int, float, bool, type(None))


def deepcopy(obj):
    """Deep copy of a wire-format object (the DeepCopy() the reference generates per type, e.g.
    api/pytorch/v1/zz_generated.deepcopy.go).  Objects here are JSON trees — dicts and lists of scalars, no
    aliasing, no cycles — so the copy needs none of copy.deepcopy's memo / dispatch machinery, which was 80 % of a
    job's submit-to-pods-created time; any other type inside still gets the general copy."""
    t = type(obj)
    if t is dict:
        return {k: (v if type(v) in _SCALARS else deepcopy(v)) for k, v in obj.items()}
    if t is list:
        return [(v if type(v) in _SCALARS else deepcopy(v)) for v in obj]
    if t in _SCALARS:
        return obj
    return copy.deepcopy(obj)


def normalize_keys(specs: Optional[Dict[str, dict]], canonical: Iterable[str]) -> None:
    """Case-insensitive replica-type keys are rewritten to their canonical spelling
    (reference api/pytorch/v1/defaults.go:78-93, api/tensorflow/v1/defaults.go:60-90)."""
    if not specs:
        return
    canon = {c.lower(): c for c in canonical}
    for key in list(specs.keys()):
        want = canon.get(key.lower())
        if want is not None and want != key:
            specs[want] = specs.pop(key)
