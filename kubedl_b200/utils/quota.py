"""Container resource arithmetic (reference pkg/util/quota/resources.go:9-35): pod request =
max(sum of containers, max of init containers), per resource name."""
from __future__ import annotations

import re
from typing import Dict, List

_SUFFIX = {"Ki": 2**10, "Mi": 2**20, "Gi": 2**30, "Ti": 2**40, "Pi": 2**50, "Ei": 2**60,
           "n": 1e-9, "u": 1e-6, "m": 1e-3, "k": 1e3, "M": 1e6, "G": 1e9, "T": 1e12, "P": 1e15, "E": 1e18}


def parse_quantity(q) -> float:
    if isinstance(q, (int, float)):
        return float(q)
    m = re.fullmatch(r"([+-]?[0-9.]+(?:[eE][+-]?[0-9]+)?)([A-Za-z]*)", str(q).strip())
    if not m:
        raise ValueError(f"bad quantity {q!r}")
    num, suf = float(m.group(1)), m.group(2)
    if suf == "":
        return num
    if suf not in _SUFFIX:
        raise ValueError(f"bad quantity suffix {q!r}")
    return num * _SUFFIX[suf]


def format_quantity(name: str, v: float) -> str:
    if name == "cpu":
        mv = round(v * 1000)
        return str(int(v)) if mv % 1000 == 0 else f"{mv}m"
    if v >= 2**30 and v % 2**30 == 0:
        return f"{int(v // 2**30)}Gi"
    if v >= 2**20 and v % 2**20 == 0:
        return f"{int(v // 2**20)}Mi"
    if v >= 2**10 and v % 2**10 == 0:
        return f"{int(v // 2**10)}Ki"
    return str(int(v)) if float(v).is_integer() else str(v)


def sum_up_containers_resources(containers: List[dict]) -> Dict[str, Dict[str, float]]:
    out = {"requests": {}, "limits": {}}
    for ctr in containers or []:
        res = ctr.get("resources") or {}
        for kind in ("requests", "limits"):
            for name, q in (res.get(kind) or {}).items():
                out[kind][name] = out[kind].get(name, 0.0) + parse_quantity(q)
    return out


def maximum_containers_resources(containers: List[dict]) -> Dict[str, Dict[str, float]]:
    out = {"requests": {}, "limits": {}}
    for ctr in containers or []:
        res = ctr.get("resources") or {}
        for kind in ("requests", "limits"):
            for name, q in (res.get(kind) or {}).items():
                out[kind][name] = max(out[kind].get(name, 0.0), parse_quantity(q))
    return out


def pod_resources(pod_spec: dict) -> Dict[str, Dict[str, float]]:
    s = sum_up_containers_resources(pod_spec.get("containers"))
    m = maximum_containers_resources(pod_spec.get("initContainers"))
    out = {"requests": {}, "limits": {}}
    for kind in ("requests", "limits"):
        for name in set(s[kind]) | set(m[kind]):
            out[kind][name] = max(s[kind].get(name, 0.0), m[kind].get(name, 0.0))
    return out


def resources_to_wire(res: Dict[str, Dict[str, float]]) -> Dict[str, Dict[str, str]]:
    return {kind: {n: format_quantity(n, v) for n, v in sorted(vals.items())} for kind, vals in res.items()}
