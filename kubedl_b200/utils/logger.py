"""Structured loggers with job / replica / pod fields (reference pkg/util/logger.go:26-80)."""
from __future__ import annotations

import logging
import os

from ..api import common as c

_root = logging.getLogger("kubedl")
if not _root.handlers:
    h = logging.StreamHandler()
    h.setFormatter(logging.Formatter("%(asctime)s %(levelname).1s %(name)s %(message)s"))
    _root.addHandler(h)
    _root.setLevel(os.environ.get("KUBEDL_LOG_LEVEL", "WARNING").upper())
    _root.propagate = False


class _Adapter(logging.LoggerAdapter):
    def process(self, msg, kwargs):
        fields = " ".join(f"{k}={v}" for k, v in self.extra.items() if v not in (None, ""))
        return (f"{msg} [{fields}]" if fields else msg), kwargs


def logger(name: str = "kubedl") -> logging.Logger:
    return logging.getLogger(name if name.startswith("kubedl") else f"kubedl.{name}")


def logger_for_job(job: dict):
    return _Adapter(_root, {"job": f"{c.namespace_of(job)}.{c.name_of(job)}", "uid": c.uid_of(job)})


def logger_for_replica(job: dict, rtype: str):
    return _Adapter(_root, {"job": f"{c.namespace_of(job)}.{c.name_of(job)}", "uid": c.uid_of(job),
                            "replica-type": rtype})


def logger_for_pod(pod: dict, kind: str = ""):
    lab = c.labels_of(pod)
    return _Adapter(_root, {"job": f"{c.namespace_of(pod)}.{lab.get(c.JOB_NAME_LABEL, '')}",
                            "pod": f"{c.namespace_of(pod)}.{c.name_of(pod)}", "kind": kind})


def logger_for_key(key: str):
    return _Adapter(_root, {"job": key})
