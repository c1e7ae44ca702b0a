"""Every manifest under examples/ is a valid object of its kind and reconciles into the pods and services it
describes (the reference ships one example per kind, example/**; these are this repository's own)."""
import glob
import os

import pytest
import yaml

from cp_helpers import Cluster
from kubedl_b200.api import crd
from kubedl_b200.api.kinds import SCHEME

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MANIFESTS = sorted(glob.glob(os.path.join(ROOT, "examples", "**", "*.yaml"), recursive=True))


def test_there_is_an_example_per_kind():
    kinds = set()
    for path in MANIFESTS:
        with open(path) as f:
            kinds |= {d["kind"] for d in yaml.safe_load_all(f) if d}
    assert {"PyTorchJob", "TFJob", "XGBoostJob", "XDLJob"} <= kinds


@pytest.mark.parametrize("path", MANIFESTS, ids=[os.path.relpath(p, ROOT) for p in MANIFESTS])
def test_example_validates_and_reconciles(path):
    with open(path) as f:
        docs = [d for d in yaml.safe_load_all(f) if d]
    jobs = [d for d in docs if SCHEME.is_workload(d.get("kind", ""))]
    assert jobs, path
    for job in jobs:
        crd.validate(job)
        cl = Cluster()
        cl.submit(job)
        specs = SCHEME.info(job["kind"]).replica_specs(cl.job(job["kind"], job["metadata"].get("namespace", "default"), job["metadata"]["name"]))
        want = sum(int(s.get("replicas", 1)) for s in specs.values())
        pods = cl.pods(job["metadata"].get("namespace", "default"))
        assert len(pods) == want, (path, len(pods), want)
        for p in pods:
            assert p["spec"]["containers"][0]["image"]
