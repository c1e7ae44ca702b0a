"""Warm GPU workers on a real device: the process that became the pod had created its CUDA context and loaded the
kernel library BEFORE the pod existed, and the pod's training runs in it unchanged (`node/zygote.py`)."""
import json
import os
import sys
import time

import pytest

pytestmark = pytest.mark.gpu


def _wait(fn, timeout=120.0):
    t0 = time.time()
    while time.time() - t0 < timeout:
        v = fn()
        if v:
            return v
        time.sleep(0.05)
    raise AssertionError("timed out")


def test_resnet_worker_in_an_adopted_warm_process(tmp_path):
    from kubedl_b200.node.zygote import ZygoteClient
    z = ZygoteClient()
    z.start()
    try:
        assert z.ready.wait(300)
        base = dict(os.environ, PYTHONPATH=os.getcwd())
        assert z.warm(0, 0, base)
        w = _wait(lambda: z.warm_workers.get(0) if z.warm_workers.get(0, {}).get("state") == "ready" else None)
        assert w["cuda"] is True and w["seconds"] > 0
        log = str(tmp_path / "pod.log")
        env = dict(base, WORLD_SIZE="1", RANK="0", KUBEDL_GPU_INDEX="0", LOCAL_RANK="0", OMP_NUM_THREADS="4",
                   KUBEDL_POD_START_TS=repr(time.time()))
        p = z.spawn([sys.executable, "-m", "kubedl_b200.worker.ddp", "--steps", "3", "--batch", "16", "--image-size", "64"],
                    env, log, gpu=0)
        assert p is not None and p.pid == w["pid"] and z.adopted == 1
        assert p.wait(300) == 0, open(log).read()
        out = [json.loads(l) for l in open(log) if l.startswith("{")][-1]
        assert out["world"] == 1 and out["loss"] == out["loss"]
        # the context was there already: no multi-second CUDA initialisation inside the pod's own clock
        assert out["startup_s"]["rendezvous"] < 1.0, out
        assert out["pod_start_to_first_step_s"] is not None and out["pod_start_to_first_step_s"] < 30
    finally:
        z.stop()
