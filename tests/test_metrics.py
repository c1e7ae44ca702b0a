

def test_per_job_launch_delay_series_are_bounded(monkeypatch):
    """The launch-delay histograms carry one series per job (name / namespace / uid labels, reference
    pkg/metrics/job_metrics.go:53-60).  The reference never drops them; here the oldest jobs' series go once more than
    KUBEDL_LAUNCH_DELAY_SERIES_MAX jobs hold one."""
    from prometheus_client import CollectorRegistry
    from kubedl_b200.plugins import metrics as M
    monkeypatch.setenv("KUBEDL_LAUNCH_DELAY_SERIES_MAX", "5")
    reg = CollectorRegistry()
    m = M.JobMetrics("PyTorchJob", None, reg)
    for i in range(12):
        job = {"metadata": {"name": f"j{i}", "namespace": "ns", "uid": f"u{i}", "creationTimestamp": "2026-01-01T00:00:00Z"}}
        pod = {"status": {"phase": "Running", "conditions": [{"type": "Ready", "status": "True",
                                                             "lastTransitionTime": "2026-01-01T00:00:02Z"}]}}
        status = {"conditions": [{"type": "Running", "status": "True"}], "startTime": "2026-01-01T00:00:01Z"}
        m.first_pod_launch_delay_seconds([pod], job, status)
        m.all_pods_launch_delay_seconds([pod], job, status)

    def series(name):
        return sorted({s.labels["name"] for fam in reg.collect() if fam.name == name for s in fam.samples
                       if s.name.endswith("_count")}, key=lambda n: int(n[1:]))
    assert series("kubedl_jobs_first_pod_launch_delay_seconds") == [f"j{i}" for i in range(7, 12)]
    assert series("kubedl_jobs_all_pods_launch_delay_seconds") == [f"j{i}" for i in range(7, 12)]
    assert len(m.launch_delays["all_pods"]) == 12            # the raw observations (bounded deque) are all there
