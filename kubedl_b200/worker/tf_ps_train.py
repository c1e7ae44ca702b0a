"""TFJob worker runtime: parameter-server training wired only by `TF_CONFIG`.

The reference launches TensorFlow images and stops there; what makes a TFJob a *distributed* job
is the `TF_CONFIG` it injects (reference controllers/tensorflow/tensorflow.go:87-141: the cluster
map of `host:port` lists per replica type plus this task's type and index).  This entrypoint is
the runtime that consumes it, in the between-graph PS/worker style the stock
`mnist_with_summaries.py` image uses (reference example/tf/tf_job_mnist.yaml:15-28):

  * every `ps` task owns one contiguous shard of the flat fp32 parameter vector (and its momentum)
    and listens on ITS OWN `host:port` entry of the cluster map — there is no global rendezvous,
    exactly like TensorFlow's gRPC servers, so workers may start, fail and be restarted by the
    operator's ExitCode policy at any time;
  * `worker` / `chief` / `master` tasks pull the shards, run forward + backward on their slice of
    the data, and push gradient slices back; the PS applies momentum SGD on arrival
    (asynchronous, the TF default) or, with `--sync_replicas`, once per step after all workers
    have reported, averaging them;
  * the chief (or master, or worker 0 when there is neither — reference
    controllers/tensorflow/status.go:60-70 uses the same rule for job success) evaluates at the
    end and writes the summaries; PS tasks serve until the operator cleans them up.

Without `TF_CONFIG` (the single-replica example) the same loop runs in-process.
Wire format: 16-byte header `!4sIQ` (op, argument, payload bytes) + raw fp32 payload.
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import socketserver
import struct
import sys
import threading
import time
from typing import Dict, List, Optional, Tuple

if __name__ == "__main__" and os.environ.get("KUBEDL_TF_TRAIN_SECONDS"):
    # lifecycle-only mode (launch-delay bench): answer before paying for `import torch`
    from kubedl_b200.worker import tf_stub
    sys.exit(tf_stub.main())

import torch

from .rendezvous import replica_identity, tf_config

HDR = struct.Struct("!4sIQ")
IN_DIM, HIDDEN, CLASSES = 784, 128, 10


# ------------------------------------------------------------------------------ model + data
def build_model(seed: int = 0) -> torch.nn.Module:
    g = torch.Generator().manual_seed(seed)
    m = torch.nn.Sequential(torch.nn.Linear(IN_DIM, HIDDEN), torch.nn.ReLU(), torch.nn.Linear(HIDDEN, CLASSES))
    with torch.no_grad():
        for p in m.parameters():
            p.copy_(torch.randn(p.shape, generator=g) * (0.05 if p.dim() > 1 else 0.0))
    return m


def flat_params(m: torch.nn.Module) -> torch.Tensor:
    return torch.cat([p.detach().reshape(-1).float().cpu() for p in m.parameters()])


def load_flat(m: torch.nn.Module, flat: torch.Tensor) -> None:
    o = 0
    with torch.no_grad():
        for p in m.parameters():
            n = p.numel()
            p.copy_(flat[o:o + n].view_as(p))
            o += n


def flat_grads(m: torch.nn.Module) -> torch.Tensor:
    return torch.cat([p.grad.detach().reshape(-1).float().cpu() for p in m.parameters()])


class Synthetic:
    """MNIST-shaped task with structure to learn: labels come from a fixed random teacher."""

    def __init__(self, worker: int, batch: int):
        # ten class prototypes; a sample is its class prototype plus unit noise
        self.protos = 0.25 * torch.randn(CLASSES, IN_DIM, generator=torch.Generator().manual_seed(1234))
        self.gen = torch.Generator().manual_seed(1000 + worker)
        self.batch = batch

    def next(self) -> Tuple[torch.Tensor, torch.Tensor]:
        y = torch.randint(0, CLASSES, (self.batch,), generator=self.gen)
        return self.protos[y] + torch.randn(self.batch, IN_DIM, generator=self.gen), y


def shard_bounds(total: int, n_shards: int) -> List[Tuple[int, int]]:
    per = (total + n_shards - 1) // n_shards
    return [(min(i * per, total), min((i + 1) * per, total)) for i in range(n_shards)]


# ------------------------------------------------------------------------------ wire helpers
def _recv_exact(sock: socket.socket, n: int) -> bytes:
    buf = bytearray()
    while len(buf) < n:
        chunk = sock.recv(min(n - len(buf), 1 << 20))
        if not chunk:
            raise ConnectionError("peer closed the connection")
        buf += chunk
    return bytes(buf)


def send_msg(sock: socket.socket, op: bytes, arg: int = 0, payload: bytes = b"") -> None:
    sock.sendall(HDR.pack(op, arg, len(payload)) + payload)


def recv_msg(sock: socket.socket) -> Tuple[bytes, int, bytes]:
    op, arg, n = HDR.unpack(_recv_exact(sock, HDR.size))
    return op, arg, _recv_exact(sock, n) if n else b""


def _tensor(payload: bytes) -> torch.Tensor:
    return torch.frombuffer(bytearray(payload), dtype=torch.float32)


# ------------------------------------------------------------------------------ parameter server
class ParameterShard:
    """One PS task's state: weights, momentum, version; async or synchronous application."""

    def __init__(self, init: torch.Tensor, lr: float, momentum: float, n_workers: int, sync: bool):
        self.w = init.clone()
        self.m = torch.zeros_like(init)
        self.lr, self.mu, self.n_workers, self.sync = lr, momentum, n_workers, sync
        self.version = 0
        self.done = 0
        self.pushes = 0
        self._acc: Optional[torch.Tensor] = None
        self._acc_n = 0
        self.cv = threading.Condition()
        self.stop = threading.Event()

    def _apply(self, g: torch.Tensor) -> None:
        self.m.mul_(self.mu).add_(g)
        self.w.add_(self.m, alpha=-self.lr)
        self.version += 1

    def push(self, g: torch.Tensor) -> int:
        with self.cv:
            self.pushes += 1
            if not self.sync:
                self._apply(g)
            else:
                self._acc = g.clone() if self._acc is None else self._acc.add_(g)
                self._acc_n += 1
                # workers that already finished no longer contribute to a round
                if self._acc_n >= max(1, self.n_workers - self.done):
                    self._apply(self._acc / self._acc_n)
                    self._acc, self._acc_n = None, 0
            self.cv.notify_all()
            return self.version

    def pull(self, min_version: int, timeout: float = 120.0) -> Tuple[int, bytes]:
        with self.cv:
            self.cv.wait_for(lambda: self.version >= min_version or self.stop.is_set(), timeout)
            return self.version, self.w.numpy().tobytes()

    # ---- checkpoint / resume: a PS pod recreated by the operator (same name, same index) picks up
    # its shard where the previous incarnation left it
    def save(self, path: str) -> None:
        with self.cv:
            state = {"w": self.w.clone(), "m": self.m.clone(), "version": self.version, "pushes": self.pushes}
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        tmp = f"{path}.tmp{os.getpid()}"
        torch.save(state, tmp)
        os.replace(tmp, path)

    def load(self, path: str) -> bool:
        if not os.path.exists(path):
            return False
        state = torch.load(path, map_location="cpu", weights_only=False)
        if state["w"].shape != self.w.shape:
            raise ValueError(f"checkpoint {path} holds {tuple(state['w'].shape)} parameters, this shard has {tuple(self.w.shape)}")
        with self.cv:
            self.w.copy_(state["w"])
            self.m.copy_(state["m"])
            self.version, self.pushes = int(state["version"]), int(state.get("pushes", 0))
            self.cv.notify_all()
        return True

    def worker_done(self) -> int:
        with self.cv:
            self.done += 1
            if self.sync and self._acc is not None and self._acc_n >= max(1, self.n_workers - self.done):
                self._apply(self._acc / self._acc_n)     # do not leave the others waiting for us
                self._acc, self._acc_n = None, 0
            self.cv.notify_all()
            return self.done


def serve_shard(shard: ParameterShard, host: str, port: int) -> socketserver.ThreadingTCPServer:
    class Handler(socketserver.BaseRequestHandler):
        def handle(self):
            sock = self.request
            sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            try:
                while not shard.stop.is_set():
                    op, arg, payload = recv_msg(sock)
                    if op == b"PULL":
                        v, data = shard.pull(arg)
                        send_msg(sock, b"OK  ", v, data)
                    elif op == b"PUSH":
                        send_msg(sock, b"OK  ", shard.push(_tensor(payload)))
                    elif op == b"DONE":
                        send_msg(sock, b"OK  ", shard.worker_done())
                    elif op == b"STAT":
                        send_msg(sock, b"OK  ", shard.version,
                                 json.dumps({"version": shard.version, "done": shard.done, "pushes": shard.pushes}).encode())
                    elif op == b"STOP":
                        shard.stop.set()
                        with shard.cv:
                            shard.cv.notify_all()
                        send_msg(sock, b"OK  ", shard.version)
                    else:
                        send_msg(sock, b"ERR ", 0, b"unknown op")
            except (ConnectionError, OSError):
                pass

    class Server(socketserver.ThreadingTCPServer):
        daemon_threads = True
        allow_reuse_address = True

    srv = Server((host, port), Handler)
    threading.Thread(target=srv.serve_forever, name="ps-shard", daemon=True).start()
    return srv


# ------------------------------------------------------------------------------ worker side
class ShardClient:
    """Connection of one worker to one PS task; reconnects while the PS pod is still starting
    (or being restarted), because nothing orders pod start-up across replica types."""

    def __init__(self, addr: str, connect_timeout: float = 120.0):
        host, _, port = addr.rpartition(":")
        self.addr = (host, int(port))
        self.connect_timeout = connect_timeout
        self.sock: Optional[socket.socket] = None

    def _connect(self) -> socket.socket:
        deadline = time.monotonic() + self.connect_timeout
        delay = 0.02
        while True:
            try:
                s = socket.create_connection(self.addr, timeout=150)
                s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
                return s
            except OSError as e:
                if time.monotonic() > deadline:
                    raise ConnectionError(f"parameter server {self.addr[0]}:{self.addr[1]} unreachable: {e}")
                time.sleep(delay)
                delay = min(delay * 2, 1.0)

    def call(self, op: bytes, arg: int = 0, payload: bytes = b"") -> Tuple[int, bytes]:
        for attempt in range(2):
            if self.sock is None:
                self.sock = self._connect()
            try:
                send_msg(self.sock, op, arg, payload)
                rop, rarg, data = recv_msg(self.sock)
                if rop != b"OK  ":
                    raise RuntimeError(f"parameter server error: {data!r}")
                return rarg, data
            except (ConnectionError, OSError):
                self.close()
                if attempt:
                    raise
        raise AssertionError("unreachable")

    def close(self) -> None:
        if self.sock is not None:
            try:
                self.sock.close()
            except OSError:
                pass
            self.sock = None


class LocalShard:
    """Same calls as ShardClient, served in-process (single-replica job: no TF_CONFIG)."""

    def __init__(self, shard: ParameterShard):
        self.shard = shard

    def call(self, op: bytes, arg: int = 0, payload: bytes = b"") -> Tuple[int, bytes]:
        if op == b"PULL":
            return self.shard.pull(arg)
        if op == b"PUSH":
            return self.shard.push(_tensor(payload)), b""
        if op == b"DONE":
            return self.shard.worker_done(), b""
        return self.shard.version, b""

    def close(self) -> None:
        pass


def train_worker(shards, bounds: List[Tuple[int, int]], worker_index: int, args, is_chief: bool,
                 device: torch.device) -> Dict[str, float]:
    model = build_model().to(device)
    data = Synthetic(worker_index, args.batch_size)
    loss_fn = torch.nn.CrossEntropyLoss()
    total = bounds[-1][1]
    flat = torch.empty(total)
    summaries = []
    first = last = None
    t0 = time.perf_counter()
    for step in range(args.max_steps):
        # synchronous mode: step s trains on version >= s (everyone's step s-1 has been applied)
        for sh, (lo, hi) in zip(shards, bounds):
            _, raw = sh.call(b"PULL", step if args.sync_replicas else 0)
            flat[lo:hi] = _tensor(raw)
        load_flat(model, flat.to(device))
        x, y = data.next()
        model.zero_grad(set_to_none=True)
        loss = loss_fn(model(x.to(device)), y.to(device))
        loss.backward()
        g = flat_grads(model)
        for sh, (lo, hi) in zip(shards, bounds):
            sh.call(b"PUSH", step, g[lo:hi].numpy().tobytes())
        last = float(loss.detach())
        first = last if first is None else first
        if is_chief and (step % args.summary_every == 0 or step == args.max_steps - 1):
            summaries.append({"step": step, "loss": last, "wall": time.perf_counter() - t0})
    for sh in shards:
        sh.call(b"DONE")
    out = {"first_loss": first, "last_loss": last, "steps": args.max_steps, "seconds": time.perf_counter() - t0}
    if is_chief:
        # evaluate what the servers hold now on a fixed held-out batch
        for sh, (lo, hi) in zip(shards, bounds):
            _, raw = sh.call(b"PULL", 0)
            flat[lo:hi] = _tensor(raw)
        load_flat(model, flat.to(device))
        ev = Synthetic(10_000, 2048)
        x, y = ev.next()
        with torch.no_grad():
            logits = model(x.to(device))
            out["eval_loss"] = float(loss_fn(logits, y.to(device)))
            out["eval_accuracy"] = float((logits.argmax(1).cpu() == y).float().mean())
        if args.log_dir:
            try:
                os.makedirs(args.log_dir, exist_ok=True)
                with open(os.path.join(args.log_dir, "summaries.jsonl"), "w") as f:
                    for s in summaries:
                        f.write(json.dumps(s) + "\n")
                    f.write(json.dumps({"eval": out}) + "\n")
            except OSError as e:      # the example mounts a hostPath that may not exist here
                print(f"summaries not written to {args.log_dir}: {e}", flush=True)
    return out


def chief_of(cfg: Optional[dict]) -> Tuple[str, int]:
    """Which task evaluates and decides the job: chief, else master, else worker 0."""
    cluster = (cfg or {}).get("cluster", {})
    for t in ("chief", "master"):
        if cluster.get(t):
            return t, 0
    return "worker", 0


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(prog="tf_ps_train")
    # flags of the stock mnist_with_summaries.py image (reference example/tf/tf_job_mnist.yaml:17-22)
    ap.add_argument("--learning_rate", type=float, default=0.05)
    ap.add_argument("--batch_size", type=int, default=100)
    ap.add_argument("--max_steps", type=int, default=int(os.environ.get("KUBEDL_TF_STEPS", "60")))
    ap.add_argument("--log_dir", default="")
    # ours
    ap.add_argument("--momentum", type=float, default=0.9)
    ap.add_argument("--sync_replicas", action="store_true", default=os.environ.get("KUBEDL_TF_SYNC", "0") == "1")
    ap.add_argument("--summary_every", type=int, default=10)
    ap.add_argument("--checkpoint_dir", default=os.environ.get("KUBEDL_TF_CHECKPOINT_DIR", ""),
                    help="ps tasks keep their shard here (ps-<index>.pt) and resume from it")
    ap.add_argument("--checkpoint_every", type=float, default=5.0, help="seconds between shard checkpoints")
    ap.add_argument("--original-command", default="", help="command line of the image this entrypoint stands in for")
    return ap


def main(argv=None) -> int:
    if os.environ.get("KUBEDL_TF_TRAIN_SECONDS"):       # lifecycle-only mode (launch-delay bench)
        from . import tf_stub
        return tf_stub.main(argv)
    args, _ = build_parser().parse_known_args(argv)
    # several replicas share a box: with torch's default of one intra-op thread per core each, six pods
    # on eight cores spend their time fighting (0.3 s/step); this model needs one thread
    torch.set_num_threads(max(1, int(os.environ.get("KUBEDL_TF_THREADS", "1"))))
    cfg = tf_config()
    total = flat_params(build_model()).numel()
    if cfg is None:
        print("TF_CONFIG absent: single local replica, in-process parameter shard", flush=True)
        shard = ParameterShard(flat_params(build_model()), args.learning_rate, args.momentum, 1, False)
        out = train_worker([LocalShard(shard)], [(0, total)], 0, args, True, _device())
        print("result " + json.dumps(out), flush=True)
        return 0 if out["last_loss"] == out["last_loss"] else 1
    task, cluster = cfg["task"], cfg["cluster"]
    ttype, tindex = task["type"].lower(), int(task["index"])
    rt, idx = replica_identity()
    if rt and (rt != ttype or idx != tindex):
        print(f"TF_CONFIG task {ttype}:{tindex} does not match the pod identity {rt}:{idx}", flush=True)
        return 2
    ps_addrs = list(cluster.get("ps") or [])
    trainers = [(t, i) for t in ("chief", "master", "worker") for i in range(len(cluster.get(t) or []))]
    n_workers = max(1, len(trainers))
    bounds = shard_bounds(total, max(1, len(ps_addrs)))
    if ttype == "ps":
        lo, hi = bounds[tindex]
        shard = ParameterShard(flat_params(build_model())[lo:hi], args.learning_rate, args.momentum,
                               n_workers, args.sync_replicas)
        port = int(ps_addrs[tindex].rpartition(":")[2])
        ckpt = os.path.join(args.checkpoint_dir, f"ps-{tindex}.pt") if args.checkpoint_dir else ""
        if ckpt and shard.load(ckpt):
            print(f"ps:{tindex} resumed from {ckpt} at version {shard.version}", flush=True)
        srv = serve_shard(shard, os.environ.get("KUBEDL_BIND_ADDR", "0.0.0.0"), port)
        print(f"ps:{tindex} serving parameters [{lo}, {hi}) of {total} on port {port} "
              f"({'sync' if args.sync_replicas else 'async'}, {n_workers} workers)", flush=True)
        # a deleted pod gets SIGTERM first (node agent and kubelet alike): stop serving, save, leave
        import signal as _signal
        _signal.signal(_signal.SIGTERM, lambda *_: shard.stop.set())
        deadline = time.monotonic() + float(os.environ.get("KUBEDL_TF_PS_SECONDS", "86400"))
        saved = shard.version
        while not shard.stop.wait(args.checkpoint_every if ckpt else 3600.0) and time.monotonic() < deadline:
            if ckpt and shard.version != saved:
                shard.save(ckpt)
                saved = shard.version
        if ckpt and shard.version != saved:
            shard.save(ckpt)
        srv.shutdown()
        print(f"ps:{tindex} stopping at version {shard.version} after {shard.pushes} pushes", flush=True)
        return 0
    if ttype == "evaluator":
        print("evaluator: nothing to evaluate against (the operator never creates this replica)", flush=True)
        return 0
    if (ttype, tindex) not in trainers:
        print(f"unknown task type {ttype!r} in TF_CONFIG", flush=True)
        return 2
    is_chief = (ttype, tindex) == chief_of(cfg)
    if ps_addrs:
        shards = [ShardClient(a) for a in ps_addrs]
    else:      # workers only (e.g. MultiWorkerMirrored-style spec): each trains its own replica
        print("no ps tasks in TF_CONFIG: training a private replica", flush=True)
        shards = [LocalShard(ParameterShard(flat_params(build_model()), args.learning_rate, args.momentum, 1, False))]
    try:
        out = train_worker(shards, bounds, trainers.index((ttype, tindex)), args, is_chief, _device())
    finally:
        for s in shards:
            s.close()
    print(f"{ttype}:{tindex} result " + json.dumps(out), flush=True)
    return int(os.environ.get("KUBEDL_FAIL_WITH", "0"))


def _device() -> torch.device:
    if torch.cuda.is_available() and os.environ.get("KUBEDL_TF_DEVICE", "") != "cpu":
        return torch.device("cuda", 0)      # gpulet exposes exactly one GPU to the pod
    return torch.device("cpu")


if __name__ == "__main__":
    from kubedl_b200.worker.exitcodes import run
    sys.exit(run(main))
