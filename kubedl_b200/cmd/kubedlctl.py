"""`python -m kubedl_b200.cmd.kubedlctl` — kubectl-style client for a running manager, plus an
embedded mode (`run`) that brings up control plane + node agent in-process, applies manifests,
waits for the jobs to finish and prints status and launch-delay metrics.

  kubedlctl apply -f job.yaml            kubedlctl get pytorchjobs -n kubedl
  kubedlctl describe tfjob mnist -n kubedl     kubedlctl delete xdljob x -n kubedl
  kubedlctl logs <pod> -n kubedl         kubedlctl run -f example.yaml --gpus 8 --timeout 600
"""
from __future__ import annotations

import argparse
import json
import sys
import time
import urllib.error
import urllib.parse
import urllib.request
from typing import List, Optional

import yaml

from ..api import common as c
from ..api.kinds import SCHEME
from ..utils import status as st

_ALIASES = {}
for _info in SCHEME.workload_kinds():
    for _n in (_info.kind.lower(), _info.plural, _info.singular):
        _ALIASES[_n] = _info.kind
_ALIASES.update({"pod": "Pod", "pods": "Pod", "svc": "Service", "service": "Service", "services": "Service",
                 "event": "Event", "events": "Event", "podgroup": "PodGroup", "podgroups": "PodGroup",
                 "node": "Node", "nodes": "Node", "no": "Node", "lease": "Lease", "leases": "Lease"})


def resolve_kind(word: str) -> str:
    k = _ALIASES.get(word.lower())
    if k is None:
        raise SystemExit(f"error: the server doesn't have a resource type \"{word}\"")
    return k


CLUSTER_SCOPED = ("Node", "Namespace")


def _path(kind: str, ns: str, name: Optional[str] = None) -> str:
    g, v, plural = SCHEME.resource(kind)
    prefix = f"/apis/{g}/{v}" if g else f"/api/{v}"
    p = f"{prefix}/{plural}" if kind in CLUSTER_SCOPED else f"{prefix}/namespaces/{ns}/{plural}"
    return p + (f"/{name}" if name else "")


def _http(server: str, method: str, path: str, body: Optional[bytes] = None):
    req = urllib.request.Request(f"http://{server}{path}", data=body, method=method)
    try:
        with urllib.request.urlopen(req, timeout=30) as r:
            data = r.read()
            ctype = r.headers.get("Content-Type", "")
    except urllib.error.HTTPError as e:
        msg = e.read().decode()
        try:
            msg = json.loads(msg).get("message", msg)
        except Exception:
            pass
        raise SystemExit(f"Error from server: {msg}")
    except urllib.error.URLError as e:
        raise SystemExit(f"cannot reach the manager at {server}: {e.reason} (start it with python -m kubedl_b200.cmd.manager)")
    return json.loads(data) if "json" in ctype else data.decode()


def _age(ts: Optional[str]) -> str:
    t = c.parse_time(ts)
    if t is None:
        return "<unknown>"
    s = int(time.time() - t)
    return f"{s}s" if s < 120 else (f"{s // 60}m" if s < 7200 else f"{s // 3600}h")


def print_table(kind: str, items: List[dict], header: bool = True) -> None:
    if SCHEME.is_workload(kind):
        # printer columns of the CRDs (reference config/crd/bases/kubeflow.org_pytorchjobs.yaml:11-32)
        rows = [("NAME", "STATE", "AGE", "FINISHED-TTL", "MAX-LIFETIME")]
        for o in items:
            sp = o.get("spec", {})
            rows.append((c.name_of(o), st.last_condition_type(o.get("status", {})) or "", _age(o["metadata"].get("creationTimestamp")),
                         str(sp.get("ttlSecondsAfterFinished", "")), str(sp.get("activeDeadlineSeconds", ""))))
    elif kind == "Pod":
        rows = [("NAME", "STATUS", "RESTARTS", "AGE", "NODE")]
        for o in items:
            cs = (o.get("status", {}).get("containerStatuses") or [{}])[0]
            rows.append((c.name_of(o), o.get("status", {}).get("phase", "Pending"), str(cs.get("restartCount", 0)),
                         _age(o["metadata"].get("creationTimestamp")), o.get("spec", {}).get("nodeName", "")))
    elif kind == "Node":
        rows = [("NAME", "STATUS", "GPUS FREE/TOTAL", "HEARTBEAT", "AGE")]
        for o in items:
            stt = o.get("status", {})
            ready = next((cd for cd in stt.get("conditions") or [] if cd.get("type") == "Ready"), {})
            rows.append((c.name_of(o), "Ready" if ready.get("status") == "True" else "NotReady",
                         f"{stt.get('allocatable', {}).get('nvidia.com/gpu', '?')}/{stt.get('capacity', {}).get('nvidia.com/gpu', '?')}",
                         _age(ready.get("lastHeartbeatTime")) + " ago", _age(o["metadata"].get("creationTimestamp"))))
    else:
        rows = [("NAME", "AGE")] + [(c.name_of(o), _age(o["metadata"].get("creationTimestamp"))) for o in items]
    w = [max(len(r[i]) for r in rows) for i in range(len(rows[0]))]
    for r in rows if header else rows[1:]:
        print("   ".join(x.ljust(w[i]) for i, x in enumerate(r)).rstrip(), flush=True)


def watch_table(server: str, kind: str, ns: str, name: Optional[str], rv: str, seconds: float) -> None:
    """`get -w`: after the table, one row per change, as `kubectl get -w` prints them (DELETED rows are marked).
    Reads the API server's chunked watch stream (apiserver/rest.py) from `rv` for `seconds`."""
    path = _path(kind, ns) + f"?watch=true&resourceVersion={rv}&timeoutSeconds={seconds:g}"
    req = urllib.request.Request(f"http://{server}{path}")
    try:
        with urllib.request.urlopen(req, timeout=seconds + 30) as r:
            for line in r:
                if not line.strip():
                    continue
                ev = json.loads(line)
                obj = ev.get("object") or {}
                if ev.get("type") == "ERROR":
                    raise SystemExit(f"watch ended: {obj.get('message', obj)}")
                if name and c.name_of(obj) != name:
                    continue
                if ev.get("type") == "DELETED":
                    print(f"{c.name_of(obj)}   <deleted>", flush=True)
                else:
                    print_table(kind, [obj], header=False)
    except urllib.error.URLError as e:
        raise SystemExit(f"cannot reach the manager at {server}: {e.reason}")
    except KeyboardInterrupt:
        pass


def _scale_to(job: dict, total: int) -> None:
    """Shrink / grow a manifest to `total` pods: singleton types (Master, Chief, Scheduler...) keep
    their one replica, the type with the most replicas absorbs the difference."""
    specs = SCHEME.info(job["kind"]).replica_specs(job)
    if not specs:
        return
    big = max(specs, key=lambda k: int(specs[k].get("replicas") or 1))
    others = sum(int(v.get("replicas") or 1) for k, v in specs.items() if k != big)
    specs[big]["replicas"] = max(total - others, 0 if others else 1)
    if specs[big]["replicas"] == 0:
        del specs[big]


def _add_env(job: dict, kv: str) -> None:
    name, _, value = kv.partition("=")
    for spec in SCHEME.info(job["kind"]).replica_specs(job).values() if SCHEME.is_workload(job.get("kind", "")) else []:
        for ctr in spec.get("template", {}).get("spec", {}).get("containers", []):
            env = [e for e in ctr.get("env") or [] if e.get("name") != name]
            env.append({"name": name, "value": value})
            ctr["env"] = env


def cmd_run(args) -> int:
    """Embedded cluster: apply, wait for every submitted job to reach a terminal condition."""
    from prometheus_client import CollectorRegistry
    from . import manager as M
    margs = M.build_parser().parse_args(["--enable-leader-election=false", "--serve", "", "--metrics-addr", "0",
                                         "--controller-metrics-addr", "0", "--workloads", "*"] + (["--gpus", str(args.gpus)] if args.gpus is not None else [])
                                        + (["--gang-scheduler-name", args.gang] if args.gang else [])
                                        + (["--prefork-workers"] if args.prefork else [])
                                        + (["--warm-gpu-workers"] if args.warm_gpus else []))
    reg = CollectorRegistry()
    op = M.build_operator(margs, metrics_registry=reg)
    M.start_operator(op, margs, metrics_registry=reg)
    jobs = []
    try:
        if (args.prefork or args.warm_gpus) and op.gpulet is not None and op.gpulet.zygote is not None:
            op.gpulet.zygote.ready.wait(300)      # a long-running agent pays this once, not per job
            if args.warm_gpus:
                t0 = time.time()
                n = op.gpulet.wait_warm(300)
                print(f"warm GPU workers: {n} ready after {time.time() - t0:.1f} s")
        for f in args.filename:
            with open(f) as fh:
                for d in yaml.safe_load_all(fh):
                    if not d:
                        continue
                    if not op.mgr.api.scheme.knows(d.get("kind", "")):
                        print(f"skipped {d.get('kind')}/{d.get('metadata', {}).get('name')} (not served on a single box)")
                        continue
                    if args.replicas and SCHEME.is_workload(d.get("kind", "")):
                        _scale_to(d, args.replicas)
                    for kv in args.env or []:
                        _add_env(d, kv)
                    o = op.mgr.api.apply(d)
                    print(f"{o['kind'].lower()}/{c.name_of(o)} applied")
                    if SCHEME.is_workload(o["kind"]):
                        jobs.append((o["kind"], c.namespace_of(o), c.name_of(o)))
        deadline = time.time() + args.timeout
        rc = 0
        for kind, ns, name in jobs:
            while time.time() < deadline:
                try:
                    status = op.mgr.api.get(kind, ns, name).get("status", {})
                except Exception:
                    status = {"conditions": [{"type": "Deleted", "status": "True"}]}
                    break
                if st.is_succeeded(status) or st.is_failed(status):
                    break
                time.sleep(0.05)
            state = st.last_condition_type(status)
            print(f"{kind.lower()}/{name}: {state}  replicaStatuses={json.dumps(status.get('replicaStatuses', {}))}")
            if state != c.JOB_SUCCEEDED:
                rc = 1
        if args.logs and op.gpulet is not None:
            import glob
            import os
            for path in sorted(glob.glob(os.path.join(op.gpulet.log_dir, "*.log"))):
                base = os.path.basename(path)[:-4]
                if not any(base.startswith(f"{ns}_{name}-") for _, ns, name in jobs):
                    continue
                with open(path, errors="replace") as fh:
                    tail = fh.read().splitlines()[-args.logs:]
                for line in tail:
                    print(f"[{base}] {line}")
        for kind, rec in op.reconcilers.items():
            for which, obs in rec.metrics.launch_delays.items():
                for key, v in obs:
                    print(f"launch-delay {which} {kind} {key}: {v * 1000:.1f} ms")
        return rc
    finally:
        op.stop()


def cmd_cluster(args) -> int:
    """The same verbs against a real Kubernetes API server (or the façade) through the client the
    operator uses — for boxes where kubectl is not installed; kubectl works just as well."""
    from ..apiserver.kube import KubeAPIServer, KubeConfig
    from ..apiserver.store import ApiError
    cfg = KubeConfig.load(master=args.server if args.token_file and not args.kubeconfig else "",
                          kubeconfig=args.kubeconfig, context=args.context, token_file=args.token_file)
    cli = KubeAPIServer(cfg)
    ns = getattr(args, "namespace", None)
    if ns == "default" and cfg.namespace:
        ns = cfg.namespace                       # the context's namespace is the default, as with kubectl
    try:
        if args.cmd == "apply":
            for f in args.filename:
                with open(f) as fh:
                    for d in yaml.safe_load_all(fh):
                        if not d:
                            continue
                        if not cli.scheme.knows(d.get("kind", "")):
                            print(f"{str(d.get('kind', '')).lower()}/{d.get('metadata', {}).get('name')} skipped (apply it with kubectl)")
                            continue
                        d.setdefault("metadata", {}).setdefault("namespace", cfg.namespace or "default")
                        o = cli.apply(d)
                        print(f"{o['kind'].lower()}/{c.name_of(o)} applied")
            return 0
        if args.cmd == "logs":
            out = cli._request("GET", cli._object_path("Pod", ns, args.pod, "log"))
            print(out.get("raw", "") if isinstance(out, dict) else out, end="")
            return 0
        kind = resolve_kind(args.kind)
        if args.cmd == "get":
            items = [cli.get(kind, ns, args.name)] if args.name else cli.list(kind, ns)
            if args.output in ("json", "yaml"):
                obj = items[0] if args.name else {"items": items}
                print(json.dumps(obj, indent=2) if args.output == "json" else yaml.safe_dump(obj, sort_keys=False))
            else:
                print_table(kind, items)
            return 0
        if args.cmd == "describe":
            obj = cli.get(kind, ns, args.name)
            print(yaml.safe_dump(obj, sort_keys=False))
            mine = cli.events_for(obj)
            if mine:
                print("Events:")
                for e in mine:
                    print(f"  {e['type']:8s} {e['reason']:28s} {e['message']}")
            return 0
        if args.cmd == "delete":
            cli.delete(kind, ns, args.name)
            print(f"{kind.lower()} \"{args.name}\" deleted")
            return 0
        return 2
    except ApiError as e:
        raise SystemExit(f"Error from server ({e.reason}): {e}")
    finally:
        cli.close()


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="kubedlctl")
    ap.add_argument("--server", "-s", default="127.0.0.1:8001")
    ap.add_argument("--kubeconfig", default="", help="talk to the Kubernetes cluster of this kubeconfig instead of a local manager")
    ap.add_argument("--context", default=None)
    ap.add_argument("--in-cluster", action="store_true", help="use $KUBECONFIG or the pod's service account")
    ap.add_argument("--token-file", default="", help="bearer token of a --server started with --token-file")
    sub = ap.add_subparsers(dest="cmd", required=True)
    a = sub.add_parser("apply"); a.add_argument("-f", "--filename", required=True, action="append")
    g = sub.add_parser("get"); g.add_argument("kind"); g.add_argument("name", nargs="?"); g.add_argument("-n", "--namespace", default="default"); g.add_argument("-o", "--output", default="")
    g.add_argument("-w", "--watch", action="store_true", help="after listing, print one row per change until interrupted")
    g.add_argument("--watch-seconds", type=float, default=3600.0, help="with -w: stop after this long")
    d = sub.add_parser("describe"); d.add_argument("kind"); d.add_argument("name"); d.add_argument("-n", "--namespace", default="default")
    r = sub.add_parser("delete"); r.add_argument("kind"); r.add_argument("name"); r.add_argument("-n", "--namespace", default="default")
    l = sub.add_parser("logs"); l.add_argument("pod"); l.add_argument("-n", "--namespace", default="default")
    l.add_argument("-f", "--follow", action="store_true", help="keep printing until the pod finishes")
    l.add_argument("--follow-seconds", type=float, default=86400.0)
    x = sub.add_parser("run"); x.add_argument("-f", "--filename", required=True, action="append")
    x.add_argument("--gpus", type=int, default=None); x.add_argument("--timeout", type=float, default=600.0); x.add_argument("--gang", default="")
    x.add_argument("--replicas", type=int, default=0, help="scale every job in the manifests to this many pods")
    x.add_argument("--env", action="append", help="NAME=VALUE added to every container of every job")
    x.add_argument("--logs", type=int, default=0, help="print the last N log lines of every pod at the end")
    x.add_argument("--prefork", action="store_true", help="fork worker pods from a launcher that has torch imported")
    x.add_argument("--warm-gpus", action="store_true",
                   help="start the job only once a worker with an initialised CUDA context waits on every GPU "
                        "(what a long-running node agent has between jobs)")
    args = ap.parse_args(argv)
    if args.cmd == "run":
        return cmd_run(args)
    if args.kubeconfig or args.in_cluster or args.token_file:
        return cmd_cluster(args)
    if args.cmd == "apply":
        for f in args.filename:
            with open(f, "rb") as fh:
                for res in _http(args.server, "POST", "/apply", fh.read()):
                    print(f"{str(res.get('kind', '')).lower()}/{res.get('name')} {res.get('result')}")
        return 0
    if args.cmd == "logs":
        text = _http(args.server, "GET", f"/logs/{args.namespace}/{args.pod}")
        print(text, end="", flush=True)
        if args.follow:       # print what is appended until the pod reaches a terminal phase (or is gone)
            deadline = time.time() + args.follow_seconds
            done = False
            while not done and time.time() < deadline:
                time.sleep(0.3)
                try:
                    phase = _http(args.server, "GET", _path("Pod", args.namespace, args.pod)).get("status", {}).get("phase")
                    done = phase in ("Succeeded", "Failed")
                except SystemExit:
                    done = True
                new = _http(args.server, "GET", f"/logs/{args.namespace}/{args.pod}")
                if len(new) > len(text):
                    print(new[len(text):], end="", flush=True)
                    text = new
        return 0
    kind = resolve_kind(args.kind)
    if args.cmd == "get":
        out = _http(args.server, "GET", _path(kind, args.namespace, args.name))
        items = [out] if args.name else out["items"]
        if args.output in ("json", "yaml"):
            obj = out if args.name else {"items": items}
            print(json.dumps(obj, indent=2) if args.output == "json" else yaml.safe_dump(obj, sort_keys=False))
        else:
            print_table(kind, items)
        if args.watch:
            rv = "0" if args.name else str(out.get("metadata", {}).get("resourceVersion", "0"))
            if rv == "0":     # a single object has no list version: follow from "now"
                rv = str(_http(args.server, "GET", _path(kind, args.namespace)).get("metadata", {}).get("resourceVersion", "0"))
            watch_table(args.server, kind, args.namespace, args.name, rv, args.watch_seconds)
        return 0
    if args.cmd == "describe":
        obj = _http(args.server, "GET", _path(kind, args.namespace, args.name))
        print(yaml.safe_dump(obj, sort_keys=False))
        if SCHEME.is_workload(kind):      # the job's pods, as `kubectl get pods -l job-name=<name>` would list them
            sel = urllib.parse.quote(f"{c.JOB_NAME_LABEL}={args.name}")
            pods = _http(args.server, "GET", _path("Pod", args.namespace) + f"?labelSelector={sel}")["items"]
            if pods:
                print("Pods:")
                print_table("Pod", sorted(pods, key=c.name_of))
                print()
        evs = _http(args.server, "GET", f"/api/v1/namespaces/{args.namespace}/events")["items"]
        mine = [e for e in evs if e.get("involvedObject", {}).get("uid") == obj["metadata"]["uid"]]
        if mine:
            print("Events:")
            for e in mine:
                print(f"  {e['type']:8s} {e['reason']:28s} {e['message']}")
        return 0
    if args.cmd == "delete":
        _http(args.server, "DELETE", _path(kind, args.namespace, args.name))
        print(f"{kind.lower()} \"{args.name}\" deleted")
        return 0
    return 2


if __name__ == "__main__":
    sys.exit(main())
