"""Shared fixtures for control-plane tests: a manager with all four controllers on a fresh
in-process API server, a fake kubelet, and an isolated Prometheus registry."""
import copy

import yaml
from prometheus_client import CollectorRegistry

from kubedl_b200.api import common as c
from kubedl_b200.controllers.registry import setup_with_manager
from kubedl_b200.engine.job_controller import JobControllerConfiguration
from kubedl_b200.node.gpulet import FakeKubelet
from kubedl_b200.plugins import gang_schedule
from kubedl_b200.runtime.manager import Manager

REF = "/root/reference"


class Cluster:
    def __init__(self, gang: str = "", workloads: str = "*", max_reconciles: int = 1):
        self.mgr = Manager()
        self.api = self.mgr.api
        self.registry = CollectorRegistry()
        gang_schedule.register_gang_schedulers(self.mgr)
        cfg = JobControllerConfiguration(enable_gang_scheduling=bool(gang), gang_scheduler_name=gang,
                                         max_concurrent_reconciles=max_reconciles)
        self.recs = setup_with_manager(self.mgr, cfg, workloads, self.registry)
        self.kubelet = FakeKubelet(self.api)

    def submit(self, job: dict) -> dict:
        out = self.api.create(copy.deepcopy(job))
        self.settle()
        return out

    def settle(self) -> int:
        return self.mgr.run_until_idle()

    def job(self, kind, ns, name) -> dict:
        return self.api.get(kind, ns, name)

    def conditions(self, kind, ns, name):
        return [(x["type"], x["status"]) for x in self.job(kind, ns, name).get("status", {}).get("conditions", [])]

    def pods(self, ns, **labels):
        return self.api.list("Pod", ns, labels or None)

    def env_of(self, pod, container=0):
        return {e["name"]: e.get("value") for e in pod["spec"]["containers"][container].get("env", [])}

    def metric(self, metric_name, **labels):
        return self.registry.get_sample_value(metric_name, labels) or 0.0


def load_example(path: str):
    with open(f"{REF}/{path}") as f:
        return [d for d in yaml.safe_load_all(f) if d]


def tmpl(container, image="img", **extra):
    return {"spec": {"containers": [{"name": container, "image": image, **extra}]}}
