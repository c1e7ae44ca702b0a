"""Workload controller registry + gate (reference controllers/controllers.go:29-45,
controllers/add_*.go).  `SETUP_WITH_MANAGER_MAP[kind] = setup(mgr, config)`; `setup_with_manager`
asks the workload gate for each kind and skips the disabled ones."""
from __future__ import annotations

from typing import Callable, Dict, List

from ..api.kinds import SCHEME
from ..engine.job_controller import JobControllerConfiguration
from ..utils.logger import logger
from ..utils.workloadgate import AUTO, is_workload_enable
from .pytorch import PytorchJobReconciler
from .tensorflow import TFJobReconciler
from .xdl import XDLJobReconciler
from .xgboost import XgboostJobReconciler

log = logger("kubedl.controllers")

SETUP_WITH_MANAGER_MAP: Dict[str, Callable] = {}


def _register(cls):
    def setup(mgr, config: JobControllerConfiguration, metrics_registry=None):
        r = cls(mgr, config, metrics_registry=metrics_registry)
        r.setup_with_manager(mgr)
        return r
    SETUP_WITH_MANAGER_MAP[cls.KIND] = setup


for _cls in (PytorchJobReconciler, TFJobReconciler, XgboostJobReconciler, XDLJobReconciler):
    _register(_cls)


def setup_with_manager(mgr, config: JobControllerConfiguration, workloads: str = AUTO,
                       metrics_registry=None) -> Dict[str, object]:
    """Returns {kind: reconciler} for every enabled workload."""
    enabled = {}
    for kind, setup in SETUP_WITH_MANAGER_MAP.items():
        if not is_workload_enable(kind, workloads, crd_installed=mgr.api.crd_installed):
            log.info("workload %s disabled by the workload gate", kind)
            continue
        enabled[kind] = setup(mgr, config, metrics_registry=metrics_registry)
        log.info("workload %s controller started", kind)
    return enabled
