"""The launch-delay benchmark itself (`kubedl_b200/bench/launch_delay.py`): the three deployment shapes it measures
produce the observations behind kubedl_jobs_{first_pod,all_pods}_launch_delay_seconds (reference
pkg/metrics/job_metrics.go:139-194) for every job they submit."""
import pytest

from kubedl_b200.bench import launch_delay as LD


@pytest.mark.parametrize("mode", ["in-process", "http", "processes"])
def test_bench_modes(mode):
    # workers live long enough for the operator to see all of them Running at once, also on a loaded box
    r = LD.run("PyTorchJob", jobs=2, replicas=3, seconds=0.5, remote=(mode == "http"), processes=(mode == "processes"))
    assert r["first_pod_ms"]["n"] == 2 and r["all_pods_ms"]["n"] == 2
    assert 0 < r["first_pod_ms"]["p50"] <= r["all_pods_ms"]["p50"] < 5000
    assert r["api_creates_per_job"] == 4          # 3 pods + the master's service
    assert r["api"].startswith("http") == (mode != "in-process")
