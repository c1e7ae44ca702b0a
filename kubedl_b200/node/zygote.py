"""Pre-forked worker launcher for the node agent.

A pod on this node agent is a Python worker process, and most of what separates "pod Running"
from "first training step" is `import torch` (4–5 s warm, tens of seconds when eight ranks page
the libraries in at once on a fresh box; `profiles/examples_through_operator_8gpu.log`: 23 s to the
first step).  The zygote is one long-lived process that has already paid for those imports and
has NOT touched CUDA; every pod whose command is `python -m <module> ...` is then an `os.fork()`
of it — copy-on-write pages, no interpreter start, no import — with the pod's own environment
(`CUDA_VISIBLE_DEVICES`, rendezvous variables) installed before the module runs, so CUDA
initialises in the child exactly as in a fresh process.  The same trick as Android's zygote and
as torch's own fork-based DataLoader workers.

Protocol, JSON lines over a UNIX stream socket (one persistent connection from the node agent):
    -> {"op": "spawn", "id": 7, "argv": ["-m", "pkg.mod", "--flag"], "env": {...}, "log": "/path", "cwd": null}
    <- {"id": 7, "pid": 4242}
    <- {"exit": 4242, "status": 0}            (later, unsolicited; negative = killed by that signal)
    -> {"op": "ping"}   <- {"pong": true, "preloaded": ["torch", ...]}
Children are session leaders (`os.setsid`), so the agent can signal their whole process group.

Warm GPU workers.  What is left after the imports are gone is CUDA itself: creating a context takes
1.5–2.5 s for one process and 10–13 s each when eight processes do it at once, which is what eight pods
of a job do (`profiles/r2_examples_through_operator_8gpu.log`: 12.7 of the 16 s to the first step).  So the
agent can ask for a worker per free GPU AHEAD of any pod: a fork that initialises CUDA on that device, loads the
kernel library and then blocks on a socket until a pod is handed to it (or the socket closes).  A pod for that
GPU is then not a fork but an adoption of the waiting worker — same pid reporting, same log / env / argv
installation — and if the worker is still initialising, the request simply waits in its socket, which is never
slower than starting a second initialisation next to it.
    -> {"op": "warm", "gpu": 3, "device": 3, "env": {"CUDA_VISIBLE_DEVICES": ...}}
    <- {"warming": 3, "pid": 4300}        then, unsolicited:  {"warmed": 3, "pid": 4300, "seconds": 2.1}
    -> {"op": "spawn", ..., "gpu": 3}     <- {"id": 7, "pid": 4300, "warm": true}
    -> {"op": "drop", "gpu": 3}           (the GPU went to a process that is not ours to pre-start)
    <- {"warm_exit": 3, "pid": 4300, "status": 1}     (a waiting worker died before any pod took it)
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import runpy
import selectors
import signal
import socket
import subprocess
import sys
import threading
import time
import traceback
from typing import Dict, List, Optional

ORPHAN_GRACE_SECONDS = float(os.environ.get("KUBEDL_ZYGOTE_ORPHAN_GRACE_SECONDS", "2"))
DEFAULT_PRELOAD = ("torch", "torch.distributed", "numpy", "kubedl_b200.worker.rendezvous")


# ------------------------------------------------------------------------------ server side
def _detach(close_fds: List[int]) -> None:
    """First thing in a forked child: drop the launcher's descriptors and handlers, lead a new session."""
    for fd in close_fds:
        try:
            os.close(fd)
        except OSError:
            pass
    for sig in (signal.SIGINT, signal.SIGTERM, signal.SIGCHLD, signal.SIGPIPE):
        signal.signal(sig, signal.SIG_DFL)
    os.setsid()
    # fork happened with INT / TERM blocked: one that arrived meanwhile is delivered now, to the
    # default action, instead of to the launcher's handler inherited for a few microseconds
    signal.pthread_sigmask(signal.SIG_UNBLOCK, {signal.SIGINT, signal.SIGTERM})


def _become_pod(req: dict) -> int:
    """Install the pod's log, environment, working directory and argv, run its module; returns the exit code."""
    log_fd = os.open(req["log"], os.O_WRONLY | os.O_CREAT | os.O_APPEND, 0o644)
    null_fd = os.open(os.devnull, os.O_RDONLY)
    os.dup2(null_fd, 0)
    os.dup2(log_fd, 1)
    os.dup2(log_fd, 2)
    os.close(log_fd)
    os.close(null_fd)
    sys.stdin = open(0, "r", closefd=False)
    sys.stdout = open(1, "w", buffering=1, closefd=False)       # line-buffered, like PYTHONUNBUFFERED logs
    sys.stderr = open(2, "w", buffering=1, closefd=False)
    os.environ.clear()
    os.environ.update(req.get("env") or {})
    for p in reversed([q for q in os.environ.get("PYTHONPATH", "").split(os.pathsep) if q]):
        if p not in sys.path:
            sys.path.insert(0, p)
    if req.get("cwd"):
        os.chdir(req["cwd"])
    _apply_thread_env()
    argv = list(req["argv"])
    assert argv[0] == "-m" and len(argv) >= 2, argv
    module, args = argv[1], argv[2:]
    sys.argv = [module] + args
    try:
        runpy.run_module(module, run_name="__main__", alter_sys=True)
        return 0
    except SystemExit as e:
        return e.code if isinstance(e.code, int) else (0 if e.code is None else 1)
    except BaseException:
        traceback.print_exc()
        return 1


def _apply_thread_env() -> None:
    """OpenMP / MKL read OMP_NUM_THREADS when they are LOADED, which happened in the launcher, without the pod's
    value: a forked pod would run its CPU ops with one thread per core of the box, and eight such pods building
    their models at once spent 2.7 s where an exec'd pod (which honours the variable) spends 0.2 s."""
    n = os.environ.get("OMP_NUM_THREADS", "")
    torch = sys.modules.get("torch")
    if torch is not None and n.isdigit() and int(n) > 0:
        try:
            torch.set_num_threads(int(n))
        except Exception:
            pass


def _exit_child(code: int) -> None:
    try:
        sys.stdout.flush()
        sys.stderr.flush()
    except Exception:
        pass
    os._exit(code & 0xFF)


def _run_child(req: dict, close_fds: List[int]) -> None:
    """In the forked child: become the pod's process.  Never returns."""
    code = 1
    try:
        _detach(close_fds)
        code = _become_pod(req)
    except BaseException:
        try:
            traceback.print_exc()
        except Exception:
            pass
        code = 1
    finally:
        _exit_child(code)


def _init_gpu(device: int) -> bool:
    """Pay for CUDA now: driver initialisation, a context on `device`, the kernel library.  False when
    this box has no GPU (the waiting / adoption protocol is the same and is what the CPU tests run)."""
    torch = sys.modules.get("torch") or importlib.import_module("torch")
    if not torch.cuda.is_available():
        return False
    torch.cuda.set_device(device % max(torch.cuda.device_count(), 1))
    torch.cuda.init()
    torch.zeros(1, device="cuda")             # the primary context and the allocator's first segment
    torch.cuda.synchronize()
    try:
        from kubedl_b200.ops import lib as L
        L.lib()
    except Exception as e:                    # a pod that needs the library reports its absence itself
        print(f"zygote: warm worker could not load the kernel library: {e}", file=sys.stderr, flush=True)
    return True


def _run_warm(req: dict, close_fds: List[int], chan: socket.socket) -> None:
    """In the forked child: initialise the GPU, then wait on `chan` for the pod to become.  Never returns."""
    code = 1
    try:
        _detach(close_fds)
        os.environ.update(req.get("env") or {})      # only what CUDA reads at initialisation; the pod's full
        t0 = time.time()                             # environment is installed at adoption
        on_gpu = _init_gpu(int(req.get("device", req.get("gpu", 0))))
        # a pod may have been handed over already, in which case the launcher has closed its end: the report
        # below then fails (quietly, not by SIGPIPE) and the request is waiting in the socket
        signal.signal(signal.SIGPIPE, signal.SIG_IGN)
        try:
            chan.sendall(json.dumps({"warmed": req.get("gpu"), "seconds": round(time.time() - t0, 3),
                                     "cuda": on_gpu}).encode() + b"\n")
        except OSError:
            pass
        signal.signal(signal.SIGPIPE, signal.SIG_DFL)
        buf = bytearray()
        while b"\n" not in buf:
            data = chan.recv(1 << 20)
            if not data:                             # launcher gone or the GPU was given away: nothing to become
                os._exit(0)
            buf += data
        chan.close()
        pod = json.loads(bytes(buf).partition(b"\n")[0])
        code = _become_pod(pod)
    except BaseException:
        try:
            traceback.print_exc()
        except Exception:
            pass
        code = 1
    finally:
        _exit_child(code)


def serve(sock_path: str, preload=DEFAULT_PRELOAD) -> int:
    loaded = []
    for name in preload:
        try:
            importlib.import_module(name)
            loaded.append(name)
        except Exception as e:      # a missing optional module must not take the launcher down
            print(f"zygote: preload of {name} failed: {e}", file=sys.stderr, flush=True)
    if "torch" in sys.modules and sys.modules["torch"].cuda.is_initialized():
        print("zygote: CUDA got initialised during preload; forked workers would not be able to use it",
              file=sys.stderr, flush=True)
        return 3
    try:
        os.unlink(sock_path)
    except FileNotFoundError:
        pass
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(sock_path)
    os.chmod(sock_path, 0o600)
    srv.listen(4)
    sel = selectors.DefaultSelector()
    sel.register(srv, selectors.EVENT_READ)
    conns: Dict[socket.socket, bytearray] = {}
    owner: Dict[int, socket.socket] = {}           # child pid -> connection that asked for it
    warm: Dict[int, dict] = {}                     # gpu -> waiting worker {pid, sock, conn, buf}
    dropped = set()                                # pids of waiting workers sent away: their exit is nobody's news
    stopping = []
    signal.signal(signal.SIGTERM, lambda *_: stopping.append(1))
    signal.signal(signal.SIGINT, lambda *_: stopping.append(1))
    print(f"zygote ready on {sock_path} (preloaded: {', '.join(loaded)})", flush=True)

    def send(conn: socket.socket, msg: dict) -> None:
        try:
            conn.sendall(json.dumps(msg).encode() + b"\n")
        except OSError:
            pass

    def forget_warm(w: dict) -> None:
        try:
            sel.unregister(w["sock"])
        except (KeyError, ValueError):
            pass
        w["sock"].close()

    def warm_readable(g: int) -> None:
        w = warm.get(g)
        if w is None:
            return
        try:
            data = w["sock"].recv(65536)
        except OSError:
            data = b""
        if not data:                               # it exited; the reaper tells the agent
            try:
                sel.unregister(w["sock"])
            except (KeyError, ValueError):
                pass
            return
        w["buf"] += data
        while b"\n" in w["buf"]:
            line, _, rest = bytes(w["buf"]).partition(b"\n")
            w["buf"][:] = rest
            try:
                msg = json.loads(line)
            except ValueError:
                continue
            if "warmed" in msg and w["conn"] in conns:
                send(w["conn"], {"warmed": g, "pid": w["pid"], "seconds": msg.get("seconds"), "cuda": msg.get("cuda")})

    def handle(conn: socket.socket, req: dict) -> None:
        if req.get("op") == "ping":
            return send(conn, {"pong": True, "preloaded": loaded, "pid": os.getpid()})
        if req.get("op") == "warm":
            g = int(req["gpu"])
            if g in warm:
                return send(conn, {"warming": g, "pid": warm[g]["pid"]})
            ours, theirs = socket.socketpair()
            fds = ([srv.fileno(), sel.fileno(), ours.fileno()] + [c.fileno() for c in conns]
                   + [w["sock"].fileno() for w in warm.values()])
            signal.pthread_sigmask(signal.SIG_BLOCK, {signal.SIGINT, signal.SIGTERM})
            try:
                pid = os.fork()
                if pid == 0:
                    _run_warm(req, fds, theirs)    # never returns
            finally:
                signal.pthread_sigmask(signal.SIG_UNBLOCK, {signal.SIGINT, signal.SIGTERM})
            theirs.close()
            warm[g] = {"pid": pid, "sock": ours, "conn": conn, "buf": bytearray()}
            sel.register(ours, selectors.EVENT_READ, ("warm", g))
            return send(conn, {"warming": g, "pid": pid})
        if req.get("op") == "drop":
            w = warm.pop(int(req["gpu"]), None)
            if w is not None:
                forget_warm(w)
                dropped.add(w["pid"])
                try:
                    os.killpg(w["pid"], signal.SIGTERM)      # the exact group forked above
                except OSError:
                    pass
            return None
        if req.get("op") == "spawn":
            w = warm.get(req["gpu"]) if req.get("gpu") is not None else None
            if w is not None:
                # hand the pod to the worker already holding a CUDA context on that GPU; if it is still
                # initialising, the request waits in its socket
                try:
                    w["sock"].sendall(json.dumps(req).encode() + b"\n")
                except OSError:
                    w = None                       # it died; the reaper reports that — fork a plain one below
                else:
                    warm.pop(req["gpu"], None)
                    forget_warm(w)
                    owner[w["pid"]] = conn
                    return send(conn, {"id": req.get("id"), "pid": w["pid"], "warm": True})
            fds = [srv.fileno(), sel.fileno()] + [c.fileno() for c in conns] + [x["sock"].fileno() for x in warm.values()]
            signal.pthread_sigmask(signal.SIG_BLOCK, {signal.SIGINT, signal.SIGTERM})
            try:
                pid = os.fork()
                if pid == 0:
                    _run_child(req, fds)           # never returns
            finally:
                signal.pthread_sigmask(signal.SIG_UNBLOCK, {signal.SIGINT, signal.SIGTERM})
            owner[pid] = conn
            return send(conn, {"id": req.get("id"), "pid": pid})
        send(conn, {"id": req.get("id"), "error": f"unknown op {req.get('op')!r}"})

    orphaned_at = None                                # when the last agent connection went away
    had_client = False
    while not stopping:
        # the launcher lives for its node agent: if the agent is gone (killed, crashed) nobody would ever hear of these
        # pods again, and waiting workers would sit on their GPUs' memory for good — leave, taking them along
        if conns:
            had_client, orphaned_at = True, None
        elif had_client:
            orphaned_at = orphaned_at or time.monotonic()
            if time.monotonic() - orphaned_at > ORPHAN_GRACE_SECONDS:
                print("zygote: the node agent is gone; shutting down", file=sys.stderr, flush=True)
                break
        for key, _ in sel.select(timeout=0.05):
            if key.data is not None:
                warm_readable(key.data[1])
                continue
            if key.fileobj is srv:
                conn, _ = srv.accept()
                conns[conn] = bytearray()
                sel.register(conn, selectors.EVENT_READ)
                continue
            conn = key.fileobj
            try:
                data = conn.recv(65536)
            except OSError:
                data = b""
            if not data:
                sel.unregister(conn)
                conns.pop(conn, None)
                conn.close()
                continue
            buf = conns[conn]
            buf += data
            while b"\n" in buf:
                line, _, rest = bytes(buf).partition(b"\n")
                buf[:] = rest
                if line.strip():
                    try:
                        handle(conn, json.loads(line))
                    except Exception as e:
                        send(conn, {"error": repr(e)})
        while True:                                   # reap
            try:
                pid, status = os.waitpid(-1, os.WNOHANG)
            except ChildProcessError:
                break
            if pid == 0:
                break
            rc = -os.WTERMSIG(status) if os.WIFSIGNALED(status) else os.WEXITSTATUS(status)
            if pid in dropped:
                dropped.discard(pid)
                continue
            for g, w in list(warm.items()):
                if w["pid"] == pid:                   # died waiting (CUDA initialisation failed, or killed)
                    warm.pop(g)
                    forget_warm(w)
                    if w["conn"] in conns:
                        send(w["conn"], {"warm_exit": g, "pid": pid, "status": rc})
                    break
            else:
                conn = owner.pop(pid, None)
                if conn is not None and conn in conns:
                    send(conn, {"exit": pid, "status": rc})
    for w in warm.values():                           # waiting workers see their socket close and leave
        w["sock"].close()
    for pid in list(owner):                           # going down: take the pods along
        try:
            os.killpg(pid, signal.SIGTERM)
        except OSError:
            pass
    srv.close()
    try:
        os.unlink(sock_path)
    except OSError:
        pass
    return 0


# ------------------------------------------------------------------------------ client side
class ZygoteProc:
    """What the node agent needs of a `subprocess.Popen`: pid, poll(), returncode, wait()."""

    def __init__(self, pid: int):
        self.pid = pid
        self.adopted = False                # True: a worker that was waiting with its GPU initialised became this pod
        self.returncode: Optional[int] = None
        self._done = threading.Event()

    def _finish(self, status: int) -> None:
        self.returncode = status
        self._done.set()

    def poll(self) -> Optional[int]:
        return self.returncode

    def wait(self, timeout: Optional[float] = None) -> Optional[int]:
        self._done.wait(timeout)
        return self.returncode


class ZygoteClient:
    """Starts the zygote in the background and forks workers from it once it is ready; until then
    (and for commands that are not `python -m module`) the caller falls back to a plain exec."""

    def __init__(self, sock_path: Optional[str] = None, preload=DEFAULT_PRELOAD, env: Optional[dict] = None):
        self.sock_path = sock_path or f"/tmp/kubedl-b200-zygote-{os.getpid()}-{id(self) & 0xFFFF:x}.sock"
        self.preload = tuple(preload)
        self.ready = threading.Event()
        self.failed = threading.Event()
        self._env = env
        self._proc: Optional[subprocess.Popen] = None
        self._sock: Optional[socket.socket] = None
        self._lock = threading.Lock()
        self._send_lock = threading.Lock()
        self._next_id = 0
        self._pending: Dict[int, dict] = {}
        self._replies: Dict[int, threading.Event] = {}
        self._children: Dict[int, ZygoteProc] = {}
        self._early_exits: Dict[int, int] = {}
        self.spawned = 0
        # warm GPU workers: gpu -> {"pid", "state": "warming" | "ready", "sig", "seconds"}
        self.warm_workers: Dict[int, dict] = {}
        self.adopted = 0                     # pods that took over a waiting worker instead of forking
        self.on_ready = []                   # callbacks run once the launcher answers (e.g. warm every free GPU)

    def _send(self, msg: dict) -> bool:
        """One request line to the launcher.  Serialised by a lock of its own: the state lock must stay free for the
        reader thread while a send blocks on a full socket buffer (the launcher may itself be blocked sending replies)."""
        sock = self._sock
        if sock is None:
            return False
        data = json.dumps(msg).encode() + b"\n"
        try:
            with self._send_lock:
                sock.sendall(data)
            return True
        except OSError:
            return False

    def start(self) -> None:
        env = dict(self._env if self._env is not None else os.environ)
        env["PYTHONPATH"] = os.pathsep.join(filter(None, [os.getcwd(), env.get("PYTHONPATH", "")]))
        env.pop("CUDA_VISIBLE_DEVICES", None)
        self._proc = subprocess.Popen([sys.executable, "-m", "kubedl_b200.node.zygote", "--socket", self.sock_path,
                                       "--preload", ",".join(self.preload)],
                                      env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, start_new_session=True)
        threading.Thread(target=self._connect, name="zygote-connect", daemon=True).start()

    def _connect(self) -> None:
        deadline = time.monotonic() + 300
        while time.monotonic() < deadline:
            if self._proc is None or self._proc.poll() is not None:
                break
            try:
                s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                s.connect(self.sock_path)
                self._sock = s
                threading.Thread(target=self._reader, name="zygote-reader", daemon=True).start()
                self.ready.set()
                for cb in list(self.on_ready):
                    try:
                        cb()
                    except Exception:
                        pass
                return
            except OSError:
                time.sleep(0.05)
        self.failed.set()

    def _reader(self) -> None:
        buf = bytearray()
        sock = self._sock
        try:
            while True:
                data = sock.recv(65536)
                if not data:
                    break
                buf += data
                while b"\n" in buf:
                    line, _, rest = bytes(buf).partition(b"\n")
                    buf[:] = rest
                    msg = json.loads(line)
                    if "exit" in msg:
                        with self._lock:
                            child = self._children.pop(msg["exit"], None)
                            if child is None:
                                self._early_exits[msg["exit"]] = msg["status"]
                        if child is not None:
                            child._finish(msg["status"])
                    elif "warming" in msg or "warmed" in msg or "warm_exit" in msg:
                        with self._lock:
                            if "warm_exit" in msg:
                                w = self.warm_workers.get(msg["warm_exit"])
                                if w is not None and w.get("pid") in (None, msg.get("pid")):
                                    self.warm_workers.pop(msg["warm_exit"], None)
                            else:
                                g = msg.get("warming", msg.get("warmed"))
                                w = self.warm_workers.get(g)
                                if w is not None:
                                    w["pid"] = msg.get("pid")
                                    if "warmed" in msg:
                                        w.update(state="ready", seconds=msg.get("seconds"), cuda=msg.get("cuda"))
                    elif "id" in msg:
                        with self._lock:
                            self._pending[msg["id"]] = msg
                            ev = self._replies.get(msg["id"])
                        if ev is not None:
                            ev.set()
        except (OSError, ValueError):
            pass
        # the zygote is gone: whatever it was parenting is gone or orphaned — report it as killed
        self.ready.clear()
        self.failed.set()
        with self._lock:
            children, self._children = list(self._children.values()), {}
        for ch in children:
            ch._finish(-signal.SIGKILL)

    @staticmethod
    def eligible(argv: List[str]) -> bool:
        return len(argv) >= 3 and argv[0] == sys.executable and argv[1] == "-m"

    @staticmethod
    def cuda_signature(env: Dict[str, str]) -> Dict[str, str]:
        """The part of an environment CUDA and the loader read when a process initialises: a waiting worker
        can only become a pod whose values for these equal the ones it was started with."""
        return {k: v for k, v in env.items() if k.startswith(("CUDA_", "NVIDIA_", "LD_"))}

    def warm(self, gpu: int, device: int, base_env: Dict[str, str], cuda_env: Optional[Dict[str, str]] = None) -> bool:
        """Ask for a worker that initialises CUDA on `device` now and waits for a pod bound to `gpu`.
        `base_env` is what a pod on that GPU would get (used for the compatibility check), `cuda_env` what has
        to be set before initialisation beyond the launcher's own environment."""
        if not self.ready.is_set():
            return False
        with self._lock:
            if gpu in self.warm_workers:
                return True
            self.warm_workers[gpu] = {"pid": None, "state": "warming", "sig": self.cuda_signature(base_env)}
        if not self._send({"op": "warm", "gpu": gpu, "device": device, "env": cuda_env or {}}):
            with self._lock:
                self.warm_workers.pop(gpu, None)
            return False
        return True

    def drop(self, gpu: int) -> None:
        """Send the worker waiting on `gpu` away (the GPU goes to a process that cannot be an adoption)."""
        with self._lock:
            if self.warm_workers.pop(gpu, None) is None:
                return
        self._send({"op": "drop", "gpu": gpu})

    def spawn(self, argv: List[str], env: Dict[str, str], log_path: str, cwd: Optional[str] = None,
              timeout: float = 10.0, gpu: Optional[int] = None) -> Optional[ZygoteProc]:
        """Fork a worker — or, with `gpu`, hand the pod to the worker already waiting on that GPU; None when
        the zygote is not (yet) usable or the command is not a module run."""
        if gpu is not None and gpu in self.warm_workers:
            w = self.warm_workers.get(gpu)
            if not self.ready.is_set() or not self.eligible(argv) or w is None or w["sig"] != self.cuda_signature(env):
                self.drop(gpu)
                gpu = None
        else:
            gpu = None
        if not self.ready.is_set() or not self.eligible(argv):
            return None
        with self._lock:
            self._next_id += 1
            rid = self._next_id
            ev = self._replies[rid] = threading.Event()
            req = {"op": "spawn", "id": rid, "argv": argv[1:], "env": env, "log": log_path, "cwd": cwd}
            if gpu is not None:
                req["gpu"] = gpu
        if not self._send(req):
            with self._lock:
                self._replies.pop(rid, None)
            return None
        if not ev.wait(timeout):
            if gpu is not None:
                self.drop(gpu)
            return None
        with self._lock:
            msg = self._pending.pop(rid, {})
            self._replies.pop(rid, None)
            pid = msg.get("pid")
            if pid is None:
                return None
            child = ZygoteProc(pid)
            if gpu is not None:
                self.warm_workers.pop(gpu, None)
                if msg.get("warm"):
                    self.adopted += 1
                    child.adopted = True
            if pid in self._early_exits:
                child._finish(self._early_exits.pop(pid))
            else:
                self._children[pid] = child
            self.spawned += 1
        return child

    def stop(self) -> None:
        if self._sock is not None:
            try:
                self._sock.close()
            except OSError:
                pass
        if self._proc is not None and self._proc.poll() is None:
            self._proc.terminate()        # the exact process we started
            try:
                self._proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self._proc.kill()
        try:
            os.unlink(self.sock_path)
        except OSError:
            pass


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="kubedl-zygote")
    ap.add_argument("--socket", required=True)
    ap.add_argument("--preload", default=",".join(DEFAULT_PRELOAD))
    a = ap.parse_args(argv)
    return serve(a.socket, tuple(p for p in a.preload.split(",") if p))


if __name__ == "__main__":
    sys.exit(main())
