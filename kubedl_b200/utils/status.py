"""Job condition state machine helpers (reference pkg/util/status.go:9-137).

Reasons and the setCondition / filterOutCondition rules are kept identical:
  * once Failed, conditions are frozen;
  * a repeat of the same (type, status, reason) is a no-op;
  * Running and Restarting are mutually exclusive;
  * adding Succeeded/Failed flips an existing Running condition to status False.
"""
from __future__ import annotations

from typing import Optional

from ..api import common as c

JOB_CREATED_REASON = "JobCreated"
JOB_SUCCEEDED_REASON = "JobSucceeded"
JOB_RUNNING_REASON = "JobRunning"
JOB_FAILED_REASON = "JobFailed"
JOB_RESTARTING_REASON = "JobRestarting"


def has_condition(status: dict, cond_type: str) -> bool:
    return any(cd.get("type") == cond_type and cd.get("status") == c.CONDITION_TRUE
               for cd in status.get("conditions") or [])


def is_succeeded(status: dict) -> bool: return has_condition(status, c.JOB_SUCCEEDED)
def is_failed(status: dict) -> bool: return has_condition(status, c.JOB_FAILED)
def is_running(status: dict) -> bool: return has_condition(status, c.JOB_RUNNING)
def is_created(status: dict) -> bool: return has_condition(status, c.JOB_CREATED)
def is_restarting(status: dict) -> bool: return has_condition(status, c.JOB_RESTARTING)


def get_condition(status: dict, cond_type: str) -> Optional[dict]:
    for cd in status.get("conditions") or []:
        if cd.get("type") == cond_type:
            return cd
    return None


def new_condition(cond_type: str, reason: str, message: str) -> dict:
    now = c.now_rfc3339()
    return {"type": cond_type, "status": c.CONDITION_TRUE, "reason": reason, "message": message,
            "lastUpdateTime": now, "lastTransitionTime": now}


def _filter_out(conditions, cond_type: str):
    out = []
    for cd in conditions or []:
        if cond_type == c.JOB_RESTARTING and cd.get("type") == c.JOB_RUNNING:
            continue
        if cond_type == c.JOB_RUNNING and cd.get("type") == c.JOB_RESTARTING:
            continue
        if cd.get("type") == cond_type:
            continue
        if cond_type in (c.JOB_FAILED, c.JOB_SUCCEEDED) and cd.get("type") == c.JOB_RUNNING:
            cd = dict(cd)
            cd["status"] = c.CONDITION_FALSE
        out.append(cd)
    return out


def set_condition(status: dict, condition: dict) -> None:
    if is_failed(status):
        return
    cur = get_condition(status, condition["type"])
    if cur is not None and cur.get("status") == condition["status"] and cur.get("reason") == condition["reason"]:
        return
    if cur is not None and cur.get("status") == condition["status"]:
        condition["lastTransitionTime"] = cur.get("lastTransitionTime")
    status["conditions"] = _filter_out(status.get("conditions"), condition["type"]) + [condition]


def update_job_conditions(status: dict, cond_type: str, reason: str, message: str) -> None:
    set_condition(status, new_condition(cond_type, reason, message))


def last_condition_type(status: dict) -> str:
    """What the CRD printer column `State` shows (.status.conditions[-1:].type)."""
    conds = status.get("conditions") or []
    return conds[-1].get("type", "") if conds else ""
