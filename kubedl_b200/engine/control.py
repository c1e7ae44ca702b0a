"""Pod / Service control and controller-ref managers.

PodControl + ServiceControl create and delete children with an ownerReference and emit the
reference's event reasons (pkg/job_controller/pod_control.go:34-47,118-138;
service_control.go:33-38,42-153).  FakeServiceControl / FakePodControl are the in-memory test
doubles (service_control.go:163-234).  The ref managers adopt matching orphans and release
children whose labels no longer match (service_ref_manager.go:81-158 and k8s
PodControllerRefManager, used at controllers/pytorch/pod.go:53-66).
"""
from __future__ import annotations

import copy  # noqa: F401
from typing import Callable, Dict, List, Optional

from ..api import common as c
from ..api.common import deepcopy as _copy
from ..apiserver.store import NotFound
from ..utils.k8sutil import match_labels

FAILED_CREATE_POD_REASON = "FailedCreatePod"
SUCCESSFUL_CREATE_POD_REASON = "SuccessfulCreatePod"
FAILED_DELETE_POD_REASON = "FailedDeletePod"
SUCCESSFUL_DELETE_POD_REASON = "SuccessfulDeletePod"
FAILED_CREATE_SERVICE_REASON = "FailedCreateService"
SUCCESSFUL_CREATE_SERVICE_REASON = "SuccessfulCreateService"
FAILED_DELETE_SERVICE_REASON = "FailedDeleteService"
SUCCESSFUL_DELETE_SERVICE_REASON = "SuccessfulDeleteService"

EVENT_NORMAL, EVENT_WARNING = "Normal", "Warning"


def validate_controller_ref(ref: Optional[dict]) -> None:
    if ref is None:
        raise ValueError("controllerRef is nil")
    if not ref.get("apiVersion"):
        raise ValueError("controllerRef has empty APIVersion")
    if not ref.get("kind"):
        raise ValueError("controllerRef has empty Kind")
    if not ref.get("controller"):
        raise ValueError("controllerRef.Controller is not set to true")
    if not ref.get("blockOwnerDeletion"):
        raise ValueError("controllerRef.BlockOwnerDeletion is not set")


def get_pod_from_template(template: dict, parent: dict, controller_ref: Optional[dict]) -> dict:
    """A Pod from a PodTemplateSpec: labels, annotations, finalizers, name and spec are copied
    (reference pod_control.go:118-138)."""
    tm = template.get("metadata") or {}
    pod = {"apiVersion": "v1", "kind": "Pod",
           "metadata": {"labels": dict(tm.get("labels") or {}), "annotations": dict(tm.get("annotations") or {}),
                        "name": tm.get("name", "")},
           "spec": _copy(template.get("spec") or {})}
    if tm.get("finalizers"):
        pod["metadata"]["finalizers"] = list(tm["finalizers"])
    if controller_ref is not None:
        pod["metadata"]["ownerReferences"] = [_copy(controller_ref)]
    return pod


class PodControl:
    def __init__(self, client, recorder):
        self.client, self.recorder = client, recorder

    def create_pods_with_controller_ref(self, namespace: str, template: dict, parent: dict,
                                        controller_ref: dict) -> dict:
        validate_controller_ref(controller_ref)
        pod = get_pod_from_template(template, parent, controller_ref)
        pod["metadata"]["namespace"] = namespace
        if not pod["metadata"]["labels"]:
            raise ValueError("unable to create pods, no labels")
        try:
            created = self.client.create(pod)
        except Exception as e:
            self.recorder.eventf(parent, EVENT_WARNING, FAILED_CREATE_POD_REASON, "Error creating: %s", e)
            raise
        self.recorder.eventf(parent, EVENT_NORMAL, SUCCESSFUL_CREATE_POD_REASON, "Created pod: %s", c.name_of(created))
        return created

    def delete_pod(self, namespace: str, name: str, parent: dict) -> None:
        try:
            self.client.delete("Pod", namespace, name)
        except NotFound:
            return
        except Exception as e:
            self.recorder.eventf(parent, EVENT_WARNING, FAILED_DELETE_POD_REASON, "Error deleting: %s", e)
            raise
        self.recorder.eventf(parent, EVENT_NORMAL, SUCCESSFUL_DELETE_POD_REASON, "Deleted pod: %s", name)

    def patch_pod(self, namespace: str, name: str, mutate: Callable[[dict], None]) -> dict:
        pod = self.client.reader().get("Pod", namespace, name)
        mutate(pod)
        pod["metadata"].pop("resourceVersion", None)
        return self.client.update(pod)


class ServiceControl:
    def __init__(self, client, recorder):
        self.client, self.recorder = client, recorder

    def create_services_with_controller_ref(self, namespace: str, service: dict, parent: dict,
                                            controller_ref: dict) -> dict:
        validate_controller_ref(controller_ref)
        if not (service.get("metadata", {}).get("labels")):
            raise ValueError("unable to create Services, no labels")
        svc = _copy(service)
        svc.setdefault("apiVersion", "v1")
        svc["kind"] = "Service"
        svc["metadata"]["namespace"] = namespace
        svc["metadata"]["ownerReferences"] = [_copy(controller_ref)]
        try:
            created = self.client.create(svc)
        except Exception as e:
            self.recorder.eventf(parent, EVENT_WARNING, FAILED_CREATE_SERVICE_REASON, "Error creating: %s", e)
            raise
        self.recorder.eventf(parent, EVENT_NORMAL, SUCCESSFUL_CREATE_SERVICE_REASON, "Created service: %s",
                             c.name_of(created))
        return created

    def delete_service(self, namespace: str, name: str, parent: dict) -> None:
        try:
            self.client.delete("Service", namespace, name)
        except NotFound:
            return
        except Exception as e:
            self.recorder.eventf(parent, EVENT_WARNING, FAILED_DELETE_SERVICE_REASON, "Error deleting: %s", e)
            raise
        self.recorder.eventf(parent, EVENT_NORMAL, SUCCESSFUL_DELETE_SERVICE_REASON, "Deleted service: %s", name)

    def patch_service(self, namespace: str, name: str, mutate: Callable[[dict], None]) -> dict:
        svc = self.client.reader().get("Service", namespace, name)
        mutate(svc)
        svc["metadata"].pop("resourceVersion", None)
        return self.client.update(svc)


class FakeServiceControl:
    """Records what would have been created / deleted / patched; optional create limit + error."""

    def __init__(self):
        self.templates: List[dict] = []
        self.controller_refs: List[dict] = []
        self.delete_service_names: List[str] = []
        self.patches: List[Callable[[dict], None]] = []
        self.err: Optional[Exception] = None
        self.create_limit = 0
        self.create_call_count = 0

    def patch_service(self, namespace, name, mutate):
        self.patches.append(mutate)
        if self.err:
            raise self.err

    def create_services_with_controller_ref(self, namespace, service, parent, controller_ref):
        self.create_call_count += 1
        if self.create_limit and self.create_call_count > self.create_limit:
            raise RuntimeError(f"not creating service, limit {self.create_limit} already reached")
        self.templates.append(_copy(service))
        self.controller_refs.append(_copy(controller_ref))
        if self.err:
            raise self.err
        return service

    def delete_service(self, namespace, name, parent):
        self.delete_service_names.append(name)
        if self.err:
            raise self.err

    def clear(self):
        self.__init__()


class FakePodControl(FakeServiceControl):
    def create_pods_with_controller_ref(self, namespace, template, parent, controller_ref):
        return self.create_services_with_controller_ref(namespace, template, parent, controller_ref)

    def delete_pod(self, namespace, name, parent):
        return self.delete_service(namespace, name, parent)

    def patch_pod(self, namespace, name, mutate):
        return self.patch_service(namespace, name, mutate)


class ControllerRefManager:
    """Claim children of `kind_obj` ("Pod" / "Service") for a controller object.

    claim(objs): keep those we own and that still match the selector; release ours that stopped
    matching; adopt matching orphans — but only if `can_adopt()` (an UNCACHED re-read proving the
    controller still exists with the same UID and is not being deleted) passes."""

    def __init__(self, control, controller: dict, selector: Dict[str, str], controller_kind: tuple,
                 can_adopt: Callable[[], None], child_kind: str):
        self.control, self.controller, self.selector = control, controller, selector
        self.controller_kind = controller_kind  # (apiVersion, kind)
        self._can_adopt = can_adopt
        self._can_adopt_err: Optional[Exception] = None
        self._can_adopt_checked = False
        self.child_kind = child_kind

    def can_adopt(self) -> None:
        if not self._can_adopt_checked:
            self._can_adopt_checked = True
            try:
                self._can_adopt()
            except Exception as e:
                self._can_adopt_err = e
        if self._can_adopt_err is not None:
            raise self._can_adopt_err

    def _patch(self, obj: dict, mutate) -> None:
        if self.child_kind == "Pod":
            self.control.patch_pod(c.namespace_of(obj), c.name_of(obj), mutate)
        else:
            self.control.patch_service(c.namespace_of(obj), c.name_of(obj), mutate)

    def adopt(self, obj: dict) -> None:
        self.can_adopt()
        ref = {"apiVersion": self.controller_kind[0], "kind": self.controller_kind[1],
               "name": c.name_of(self.controller), "uid": c.uid_of(self.controller),
               "controller": True, "blockOwnerDeletion": True}

        def mutate(o):
            o["metadata"].setdefault("ownerReferences", []).append(ref)
        self._patch(obj, mutate)

    def release(self, obj: dict) -> None:
        uid = c.uid_of(self.controller)

        def mutate(o):
            o["metadata"]["ownerReferences"] = [r for r in o["metadata"].get("ownerReferences") or []
                                                if r.get("uid") != uid]
        try:
            self._patch(obj, mutate)
        except NotFound:
            pass

    def claim_object(self, obj: dict) -> bool:
        ref = c.controller_ref(obj)
        matches = match_labels(obj, self.selector)
        if ref is not None:
            if ref.get("uid") != c.uid_of(self.controller):
                return False  # owned by someone else
            if matches:
                return True
            if self.controller.get("metadata", {}).get("deletionTimestamp"):
                return False
            self.release(obj)
            return False
        # orphan
        if self.controller.get("metadata", {}).get("deletionTimestamp") or not matches:
            return False
        if obj.get("metadata", {}).get("deletionTimestamp"):
            return False
        try:
            self.adopt(obj)
        except NotFound:
            return False
        return True

    def claim(self, objs: List[dict]) -> List[dict]:
        claimed, errs = [], []
        for o in objs:
            try:
                if self.claim_object(o):
                    claimed.append(o)
            except Exception as e:
                errs.append(e)
        if errs:
            raise errs[0]
        return claimed
