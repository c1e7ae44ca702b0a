"""Code sync plugin: inject a git-sync init container into every replica of a job that carries
the `kubedl.io/git-sync-config` annotation (reference pkg/code_sync/sync_handler.go:15-73,
git_sync_handler.go:17-152; user guide reference docs/sync_code.md).

The handler contract is `init_container(opts_json, mount_volume) -> (container, dest_path)`; git
is the only implementation upstream.  On the single-box node agent the init container is run
by gpulet as a local `git clone` (node/gpulet.py), so the feature works end to end.
"""
from __future__ import annotations

import copy
import json
import posixpath
from typing import Dict, Protocol, Tuple

from ..api import common as c

DEFAULT_CODE_ROOT_PATH = "/code"
DEFAULT_GIT_SYNC_IMAGE = "kubedl/git-sync:v1"
GIT_SYNC_CONTAINER_NAME = "git-sync-code"
GIT_SYNC_VOLUME_NAME = "git-sync"


class CodeSyncHandler(Protocol):
    def init_container(self, opts_config: str, mount_volume: dict) -> Tuple[dict, str]: ...


class GitSyncHandler:
    def init_container(self, opts_config: str, mount_volume: dict) -> Tuple[dict, str]:
        opts = json.loads(opts_config)
        if not isinstance(opts, dict) or not opts.get("source"):
            raise ValueError("git-sync-config needs a `source` repository")
        root = opts.get("rootPath") or DEFAULT_CODE_ROOT_PATH
        dest = opts.get("destPath")
        if not dest:
            dest = opts["source"].strip("/").split("/")[-1]
            if dest.endswith(".git"):
                dest = dest[:-4]
        image = opts.get("image") or DEFAULT_GIT_SYNC_IMAGE
        max_failures = opts.get("maxFailures") or 3
        envs = list(opts.get("envs") or [])

        def add(name, value):
            envs.append({"name": name, "value": str(value)})
        add("GIT_SYNC_REPO", opts["source"])
        add("GIT_SYNC_ONE_TIME", "true")
        if max_failures >= 0:
            add("GIT_SYNC_MAX_SYNC_FAILURES", max_failures)
        if opts.get("branch"):
            add("GIT_SYNC_BRANCH", opts["branch"])
        if opts.get("revision"):
            add("GIT_SYNC_REV", opts["revision"])
        if opts.get("depth"):
            add("GIT_SYNC_DEPTH", opts["depth"])
        add("GIT_SYNC_ROOT", root)
        add("GIT_SYNC_DEST", dest)
        if opts.get("ssh"):
            add("GIT_SYNC_SSH", "true")
            if opts.get("sshFile"):
                add("GIT_SSH_KEY_FILE", opts["sshFile"])
        if opts.get("user"):
            add("GIT_SYNC_USERNAME", opts["user"])
        if opts.get("password"):
            add("GIT_SYNC_PASSWORD", opts["password"])
        container = {"name": GIT_SYNC_CONTAINER_NAME, "image": image, "env": envs,
                     "imagePullPolicy": "IfNotPresent",
                     "volumeMounts": [{"name": mount_volume["name"], "readOnly": False, "mountPath": root}]}
        return container, dest


def inject_code_sync_init_container(opts_config: str, handler: CodeSyncHandler, specs: Dict[str, dict],
                                    mount_volume: dict) -> None:
    init_container, dest = handler.init_container(opts_config, mount_volume)
    for spec in specs.values():
        pod_spec = spec.setdefault("template", {}).setdefault("spec", {})
        containers = pod_spec.get("containers") or []
        ic = copy.deepcopy(init_container)
        if containers:
            ic["resources"] = copy.deepcopy(containers[0].get("resources") or {})
        pod_spec.setdefault("initContainers", [])
        if any(x.get("name") == ic["name"] for x in pod_spec["initContainers"]):
            continue  # idempotent: reconcile re-enters with the same in-memory specs
        pod_spec["initContainers"].append(ic)
        pod_spec.setdefault("volumes", []).append(copy.deepcopy(mount_volume))
        for ctr in containers:
            ctr.setdefault("volumeMounts", []).append({
                "name": mount_volume["name"], "readOnly": False,
                "mountPath": posixpath.join(ctr.get("workingDir") or "", dest), "subPath": dest})


def inject_code_sync_init_containers(job: dict, specs: Dict[str, dict]) -> None:
    cfg = c.annotations_of(job).get(c.ANNOTATION_GIT_SYNC_CONFIG)
    if cfg is not None:
        inject_code_sync_init_container(cfg, GitSyncHandler(), specs,
                                        {"name": GIT_SYNC_VOLUME_NAME, "emptyDir": {}})
