"""Local emulation of the `kubedl/git-sync` init container (reference docs/sync_code.md,
pkg/code_sync/git_sync_handler.go): clone GIT_SYNC_REPO (branch / rev / depth honoured) into
$GIT_SYNC_ROOT/$GIT_SYNC_DEST, once."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys


def main() -> int:
    repo = os.environ.get("GIT_SYNC_REPO")
    if not repo:
        print("GIT_SYNC_REPO not set", file=sys.stderr)
        return 2
    root = os.environ.get("GIT_SYNC_ROOT", "/code")
    if not os.access(os.path.dirname(root) or "/", os.W_OK):
        root = os.path.join("/tmp/kubedl-b200/code", root.strip("/"))
    dest = os.path.join(root, os.environ.get("GIT_SYNC_DEST", "code"))
    os.makedirs(root, exist_ok=True)
    if os.path.exists(dest):
        shutil.rmtree(dest)
    cmd = ["git", "clone"]
    if os.environ.get("GIT_SYNC_DEPTH"):
        cmd += ["--depth", os.environ["GIT_SYNC_DEPTH"]]
    if os.environ.get("GIT_SYNC_BRANCH"):
        cmd += ["--branch", os.environ["GIT_SYNC_BRANCH"]]
    cmd += [repo, dest]
    tries = int(os.environ.get("GIT_SYNC_MAX_SYNC_FAILURES", "3")) + 1
    for _ in range(tries):
        if subprocess.call(cmd) == 0:
            rev = os.environ.get("GIT_SYNC_REV")
            if rev and subprocess.call(["git", "-C", dest, "checkout", rev]) != 0:
                return 1
            return 0
    return 1


if __name__ == "__main__":
    sys.exit(main())
