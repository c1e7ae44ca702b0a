"""Tenancy info carried in the `kubedl.io/tenancy` annotation
(reference pkg/util/tenancy/tenancy.go:25-43)."""
from __future__ import annotations

import json
from dataclasses import dataclass
from typing import Optional

from ..api import common as c


@dataclass
class Tenancy:
    tenant: str = ""
    user: str = ""
    idc: str = ""
    region: str = ""


def get_tenancy(obj: dict) -> Optional[Tenancy]:
    """Returns None when the annotation is absent; raises ValueError on malformed JSON."""
    raw = c.annotations_of(obj).get(c.ANNOTATION_TENANCY_INFO)
    if raw is None:
        return None
    try:
        d = json.loads(raw)
    except json.JSONDecodeError as e:
        raise ValueError(f"malformed {c.ANNOTATION_TENANCY_INFO} annotation: {e}") from e
    return Tenancy(tenant=d.get("tenant", ""), user=d.get("user", ""), idc=d.get("idc", ""),
                   region=d.get("region", ""))
