"""`python -m kubedl_b200.cmd.apiserver` — the API server on its own.

The all-in-one manager keeps cluster state in its own process.  For the reference's deployment
shape — state in an API server, the operator as one client of it, node agents as others
(reference config/manager/all_in_one.yaml) — run this, then point
`python -m kubedl_b200.cmd.manager --master URL` and `python -m kubedl_b200.cmd.gpulet --master URL`
at it.  It serves the Kubernetes wire protocol (`apiserver/rest.py`): `kubedlctl --server URL` and
`kubedlctl --kubeconfig` work against it (kubectl's discovery / list / watch calls are the same ones, but
kubectl is not in this image and has not been tried).

State is in memory; `--state FILE` loads a snapshot at start and writes one at exit (and every
`--snapshot-interval` seconds), which is enough for an operator restart test, not a database.
"""
from __future__ import annotations

import argparse
import json
import os
import signal
import sys
import threading

from ..api.kinds import SCHEME
from ..apiserver import rest
from ..apiserver.store import APIServer
from ..utils.logger import logger

log = logger("kubedl.apiserver")


def save_state(api: APIServer, path: str) -> int:
    objs = []
    for kind in api.kinds_present():
        if kind != "Event":
            objs.extend(api.list(kind))
    tmp = path + ".tmp"
    with open(tmp, "w") as f:
        json.dump({"resourceVersion": api.current_rv(), "items": objs}, f)
    os.replace(tmp, path)
    return len(objs)


def load_state(api: APIServer, path: str) -> int:
    with open(path) as f:
        snap = json.load(f)
    for o in snap.get("items", []):
        o.get("metadata", {}).pop("resourceVersion", None)   # uid and creationTimestamp survive
        api.create(o)
    return len(snap.get("items", []))


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="kubedl-apiserver")
    ap.add_argument("--serve", default="127.0.0.1:8001", help="listen address")
    ap.add_argument("--crds", default="*", help="comma list of workload kinds to serve (`*` = all)")
    ap.add_argument("--state", default="", help="snapshot file to load at start and save at exit")
    ap.add_argument("--token-file", default="", help="require `Authorization: Bearer <contents>`; the file is created "
                    "with a random token (mode 0600) if it does not exist.  Use it whenever --serve is not loopback")
    ap.add_argument("--snapshot-interval", type=float, default=0.0)
    args = ap.parse_args(argv)
    from .manager import low_latency_threads
    low_latency_threads()
    api = APIServer(install_all_crds=False)
    wanted = [k.strip().lower() for k in args.crds.split(",") if k.strip()]
    for info in SCHEME.workload_kinds():
        if "*" in wanted or info.kind.lower() in wanted or info.plural in wanted:
            api.install_crd(info.kind)
    if args.state and os.path.exists(args.state):
        log.warning("loaded %d objects from %s", load_state(api, args.state), args.state)
    host, _, port = args.serve.rpartition(":")
    token = ""
    if args.token_file:
        if not os.path.exists(args.token_file):
            import secrets
            fd = os.open(args.token_file, os.O_WRONLY | os.O_CREAT | os.O_EXCL, 0o600)
            with os.fdopen(fd, "w") as f:
                f.write(secrets.token_urlsafe(32))
        with open(args.token_file) as f:
            token = f.read().strip()
        if not token:
            # an empty token means "authentication off" further down: never fall into that silently
            print(f"kubedl-b200 apiserver: --token-file {args.token_file} exists but holds no token; "
                  "delete it (a random one is generated) or write a token into it", file=sys.stderr, flush=True)
            return 2
    elif (host or "127.0.0.1") not in ("127.0.0.1", "localhost", "::1"):
        log.warning("listening on %s without --token-file: anyone who can reach the port owns the cluster", host)
    srv = rest.serve(api, host or "127.0.0.1", int(port), token=token)
    print(f"kubedl-b200 apiserver listening on {srv.url}", flush=True)
    stop = threading.Event()
    for sig in (signal.SIGINT, signal.SIGTERM):
        signal.signal(sig, lambda *_: stop.set())
    while not stop.wait(args.snapshot_interval or None):
        if args.state:
            save_state(api, args.state)
    if args.state:
        save_state(api, args.state)
    srv.shutdown()
    return 0


if __name__ == "__main__":
    sys.exit(main())
