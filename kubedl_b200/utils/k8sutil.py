"""Pod filters and replica sums (reference pkg/util/k8sutil/k8sutil.go:95-160)."""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

from ..api import common as c


def phase_of(pod: dict) -> str:
    return (pod.get("status") or {}).get("phase", "") or c.POD_PENDING


def is_pod_active(pod: dict) -> bool:
    return phase_of(pod) not in (c.POD_SUCCEEDED, c.POD_FAILED) and \
        not pod.get("metadata", {}).get("deletionTimestamp")


def filter_active_pods(pods: List[dict]) -> List[dict]:
    return [p for p in pods if is_pod_active(p)]


def filter_pod_count(pods: List[dict], phase: str) -> int:
    return sum(1 for p in pods if phase_of(p) == phase)


def get_total_replicas(specs: Dict[str, dict]) -> int:
    return c.total_replicas(specs)


def get_total_failed_replicas(replica_statuses: Optional[Dict[str, dict]]) -> int:
    return sum(int(rs.get("failed", 0)) for rs in (replica_statuses or {}).values())


def get_total_active_replicas(replica_statuses: Optional[Dict[str, dict]]) -> int:
    return sum(int(rs.get("active", 0)) for rs in (replica_statuses or {}).values())


def resolve_dependent_owner(obj: dict) -> Tuple[str, str]:
    ref = c.controller_ref(obj)
    return (ref.get("uid", ""), ref.get("name", "")) if ref else ("", "")


def get_replica_type(pod: dict) -> Tuple[str, bool]:
    lab = c.labels_of(pod)
    return lab.get(c.REPLICA_TYPE_LABEL, ""), c.REPLICA_TYPE_LABEL in lab


def match_labels(obj: dict, selector: Dict[str, str]) -> bool:
    lab = c.labels_of(obj)
    return all(lab.get(k) == v for k, v in (selector or {}).items())
