"""TFJob stand-in for the stock TensorFlow image (BASELINE.json config #4: "TF_CONFIG injection
... lifecycle/metrics parity with stock TF image").  TensorFlow is not installed here, so this
entrypoint validates what the operator must provide — a well-formed TF_CONFIG whose task entry
names this replica (absent for a single-replica job, reference tfjob_controller.go:224-245) —
then behaves like a short training run: PS replicas serve until killed, everyone else exits 0."""
from __future__ import annotations

import os
import sys
import time

from .rendezvous import replica_identity, tf_config, tf_rank_world


def main(argv=None) -> int:
    cfg = tf_config()
    rt, idx = replica_identity()
    if cfg is None:
        print("TF_CONFIG absent: running as a single local replica", flush=True)
    else:
        assert cfg.get("environment") == "cloud", cfg
        task = cfg["task"]
        assert task["type"] in cfg["cluster"] or task["type"] == "evaluator", cfg
        if task["type"] in cfg["cluster"]:
            assert 0 <= int(task["index"]) < len(cfg["cluster"][task["type"]]), cfg
        if rt:
            assert task["type"] == rt and int(task["index"]) == idx, (task, rt, idx)
        rank, world = tf_rank_world(cfg) if task["type"] in cfg["cluster"] else (0, 1)
        print(f"TF_CONFIG ok: task {task['type']}:{task['index']} rank {rank}/{world}", flush=True)
        if task["type"] == "ps":
            time.sleep(float(os.environ.get("KUBEDL_TF_PS_SECONDS", "3600")))
            return 0
    time.sleep(float(os.environ.get("KUBEDL_TF_TRAIN_SECONDS", "0.2")))
    return int(os.environ.get("KUBEDL_FAIL_WITH", "0"))


if __name__ == "__main__":
    sys.exit(main())
