"""Workload gate: which kinds get a controller (reference pkg/util/workloadgate/
workload_gate.go:26-112).

`--workloads` / env `WORKLOADS_ENABLE` (env wins):
  auto            enable a kind iff its CRD is served by the API server (discovery)
  *               enable everything
  A,B,-C          enable A and B, disable C; `*,-C` = all but C
`KUBEDL_CI=true` skips discovery (reference workload_gate.go:108-112).
"""
from __future__ import annotations

import os
from typing import Callable, Dict, Optional, Tuple

ENV_WORKLOADS_ENABLE = "WORKLOADS_ENABLE"
AUTO = "auto"


def parse_workloads_enabled(spec: str) -> Tuple[Dict[str, bool], bool]:
    """-> ({kind: enabled}, enable_all).  `-*` is ignored, like the reference parser
    (workload_gate.go:63-88; cases from workload_gate_test.go:8-64)."""
    known: Dict[str, bool] = {}
    enable_all = False
    for raw in spec.split(","):
        w = raw.strip()
        enable = True
        if w.startswith("-"):
            enable = False
            w = w[1:]
        if w == "*":
            if enable:
                enable_all = True
            continue
        if not w:
            continue
        known[w] = enable
    return known, enable_all


def is_workload_enable(kind: str, workloads_flag: str = AUTO,
                       crd_installed: Optional[Callable[[str], bool]] = None) -> bool:
    """Env overrides the flag (docs/startup_flags.md).  Two reference quirks are fixed rather than
    kept: an `auto` flag no longer shadows the env setting, and `-Kind` really disables the kind
    (the reference tests map *presence*, so `-foo` enabled foo: workload_gate.go:57-60)."""
    spec = os.environ.get(ENV_WORKLOADS_ENABLE) or workloads_flag or AUTO
    if spec == AUTO:
        if os.environ.get("KUBEDL_CI", "").lower() == "true" or crd_installed is None:
            return True
        return bool(crd_installed(kind))
    known, enable_all = parse_workloads_enabled(spec)
    lowered = {k.lower(): v for k, v in known.items()}
    if kind.lower() in lowered:
        return lowered[kind.lower()]
    return enable_all
