"""CustomResourceDefinition manifests generated from the kind registry.

Same shape as the reference's controller-gen output (reference
config/crd/bases/kubeflow.org_pytorchjobs.yaml:1-160): namespaced, status subresource, printer
columns State / Age / Finished-TTL / Max-Lifetime, a validation schema that requires the
replica-spec map under .spec and `conditions` + `replicaStatuses` under .status.
`python -m kubedl_b200.api.crd --out config/crd/bases` rewrites the files (the `make manifests`
step of reference Makefile:47-48).
"""
from __future__ import annotations

import argparse
import copy
import os
from typing import Dict

import yaml

from . import common as c
from .kinds import SCHEME, KindInfo, XDL_KIND

PRINTER_COLUMNS = [
    {"JSONPath": ".status.conditions[-1:].type", "name": "State", "type": "string"},
    {"JSONPath": ".metadata.creationTimestamp", "name": "Age", "type": "date"},
    {"JSONPath": ".spec.ttlSecondsAfterFinished", "name": "Finished-TTL", "type": "integer"},
    {"JSONPath": ".spec.activeDeadlineSeconds", "name": "Max-Lifetime", "type": "integer"},
]


def _replica_spec_schema() -> dict:
    return {"type": "object", "properties": {
        "replicas": {"type": "integer", "format": "int32"},
        "restartPolicy": {"type": "string", "enum": list(c.RESTART_POLICIES)},
        "template": {"type": "object", "description": "core/v1 PodTemplateSpec"}}}


def _status_schema() -> dict:
    cond = {"type": "object", "required": ["status", "type"], "properties": {
        "type": {"type": "string", "enum": list(c.CONDITION_TYPES)}, "status": {"type": "string"},
        "reason": {"type": "string"}, "message": {"type": "string"},
        "lastUpdateTime": {"type": "string", "format": "date-time"},
        "lastTransitionTime": {"type": "string", "format": "date-time"}}}
    rs = {"type": "object", "properties": {k: {"type": "integer", "format": "int32"} for k in ("active", "succeeded", "failed")}}
    return {"type": "object", "required": ["conditions", "replicaStatuses"], "properties": {
        "conditions": {"type": "array", "items": cond},
        "replicaStatuses": {"type": "object", "additionalProperties": rs},
        "startTime": {"type": "string", "format": "date-time"},
        "completionTime": {"type": "string", "format": "date-time"},
        "lastReconcileTime": {"type": "string", "format": "date-time"}}}


def crd_for(info: KindInfo) -> dict:
    spec_props: Dict[str, dict] = {
        "cleanPodPolicy": {"type": "string", "enum": list(c.CLEAN_POD_POLICIES)},
        "ttlSecondsAfterFinished": {"type": "integer", "format": "int32"},
        "activeDeadlineSeconds": {"type": "integer", "format": "int64"},
        "backoffLimit": {"type": "integer", "format": "int32"},
        "schedulingPolicy": {"type": "object", "properties": {"minAvailable": {"type": "integer", "format": "int32"}}},
        info.replica_specs_key: {"type": "object", "additionalProperties": _replica_spec_schema(),
                                 "description": "map of replica type (%s) to ReplicaSpec" % ", ".join(info.replica_types)},
    }
    if info.kind == XDL_KIND:
        spec_props["minFinishWorkNum"] = {"type": "integer", "format": "int32"}
        spec_props["minFinishWorkRate"] = {"type": "integer", "format": "int32"}
    return {
        "apiVersion": "apiextensions.k8s.io/v1beta1", "kind": "CustomResourceDefinition",
        "metadata": {"name": info.crd_name, "annotations": {"kubedl-b200/generated-by": "kubedl_b200.api.crd"}},
        "spec": {
            "group": info.group, "version": info.version, "scope": "Namespaced",
            "names": {"kind": info.kind, "listKind": info.kind + "List", "plural": info.plural, "singular": info.singular},
            "additionalPrinterColumns": PRINTER_COLUMNS,
            "subresources": {"status": {}},
            "validation": {"openAPIV3Schema": {"type": "object", "properties": {
                "apiVersion": {"type": "string"}, "kind": {"type": "string"}, "metadata": {"type": "object"},
                "spec": {"type": "object", "required": [info.replica_specs_key], "properties": spec_props},
                "status": _status_schema()}}},
            "versions": [{"name": info.version, "served": True, "storage": True}],
        },
    }


def crd_for_v1(info: KindInfo) -> dict:
    """The same CRD as `apiextensions.k8s.io/v1` (clusters from 1.22 on no longer serve the
    v1beta1 form the reference ships).  v1 wants a structural schema per version: every object
    node is typed, and free-form parts (pod templates, metadata of embedded objects) say so."""
    beta = crd_for(info)
    schema = copy.deepcopy(beta["spec"]["validation"]["openAPIV3Schema"])

    def structural(node: dict) -> None:
        if node.get("type") == "object" and "properties" not in node and "additionalProperties" not in node:
            node["x-kubernetes-preserve-unknown-fields"] = True
        for child in (node.get("properties") or {}).values():
            structural(child)
        if isinstance(node.get("additionalProperties"), dict):
            structural(node["additionalProperties"])
        if isinstance(node.get("items"), dict):
            structural(node["items"])
    structural(schema)
    cols = [{"jsonPath": col["JSONPath"], "name": col["name"], "type": col["type"]} for col in PRINTER_COLUMNS]
    return {
        "apiVersion": "apiextensions.k8s.io/v1", "kind": "CustomResourceDefinition",
        "metadata": copy.deepcopy(beta["metadata"]),
        "spec": {"group": info.group, "scope": "Namespaced", "names": copy.deepcopy(beta["spec"]["names"]),
                 "versions": [{"name": info.version, "served": True, "storage": True,
                               "additionalPrinterColumns": cols, "subresources": {"status": {}},
                               "schema": {"openAPIV3Schema": schema}}]},
    }


def filename_for(info: KindInfo) -> str:
    return f"{info.group}_{info.plural}.yaml"


def write_all(out_dir: str) -> list:
    os.makedirs(out_dir, exist_ok=True)
    paths = []
    for info in SCHEME.workload_kinds():
        p = os.path.join(out_dir, filename_for(info))
        with open(p, "w") as f:
            f.write("# generated by `python -m kubedl_b200.api.crd`; do not edit\n---\n")
            yaml.safe_dump(crd_for(info), f, sort_keys=False)
        paths.append(p)
    return paths


def write_all_v1(out_dir: str) -> list:
    os.makedirs(out_dir, exist_ok=True)
    paths = []
    for info in SCHEME.workload_kinds():
        p = os.path.join(out_dir, filename_for(info))
        with open(p, "w") as f:
            f.write("# generated by `python -m kubedl_b200.api.crd`; do not edit\n---\n")
            yaml.safe_dump(crd_for_v1(info), f, sort_keys=False)
        paths.append(p)
    return paths


def all_in_one(config_dir: str = "config") -> str:
    """One file to `kubectl apply`: namespaces, v1 CRDs, RBAC, operator Deployment — what the
    reference ships as config/manager/all_in_one.yaml (namespaces `kubedl` and `kubedl-system`,
    ClusterRole + binding, Deployment; reference all_in_one.yaml:1-328)."""
    parts = ["# generated by `python -m kubedl_b200.api.crd`; do not edit\n"
             "# kubectl apply -f config/manager/all_in_one.yaml\n"]
    docs = [{"apiVersion": "v1", "kind": "Namespace", "metadata": {"name": "kubedl"}}]
    docs += [crd_for_v1(info) for info in SCHEME.workload_kinds()]
    for rel in ("manager/in_cluster.yaml", "rbac/role.yaml", "rbac/role_binding.yaml",
                "rbac/leader_election_role.yaml", "rbac/leader_election_role_binding.yaml"):
        with open(os.path.join(config_dir, rel)) as f:
            docs += [d for d in yaml.safe_load_all(f) if d]
    order = {"Namespace": 0, "CustomResourceDefinition": 1, "ServiceAccount": 2, "ClusterRole": 3, "Role": 3,
             "ClusterRoleBinding": 4, "RoleBinding": 4, "Deployment": 5}
    docs.sort(key=lambda d: order.get(d["kind"], 9))
    return parts[0] + "---\n".join(yaml.safe_dump(d, sort_keys=False) for d in docs)


_VALIDATORS: dict = {}


def _validator(kind: str):
    """One compiled validator per kind.  `jsonschema.validate(doc, schema)` re-generates nothing here but re-checks
    the schema against its metaschema on EVERY call, which cost ~10 ms per job create — more than the reconcile
    that follows; the schema of a kind never changes inside a process."""
    info = SCHEME.info(kind)
    hit = _VALIDATORS.get(kind)
    if hit is None or hit[0] is not info:         # a kind registered again (custom workloads, tests) gets a new validator
        import jsonschema
        schema = crd_for(info)["spec"]["validation"]["openAPIV3Schema"]
        cls = jsonschema.validators.validator_for(schema)
        cls.check_schema(schema)
        hit = _VALIDATORS[kind] = (info, cls(schema))
    return hit[1]


def validate(obj: dict) -> None:
    """Minimal structural validation of a job against the CRD schema (what the API server's
    OpenAPI validation would reject)."""
    import jsonschema
    doc = {k: v for k, v in obj.items() if k != "status"} if "status" not in obj else obj
    err = jsonschema.exceptions.best_match(_validator(obj.get("kind", "")).iter_errors(doc))
    if err is not None:
        raise err


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="config/crd/bases")
    ap.add_argument("--out-v1", default="config/crd/v1")
    a = ap.parse_args()
    for p in write_all(a.out) + write_all_v1(a.out_v1):
        print(p)
    with open("config/manager/all_in_one.yaml", "w") as f:
        f.write(all_in_one())
    print("config/manager/all_in_one.yaml")
