"""Turn worker failures into the exit codes the operator's `restartPolicy: ExitCode` understands.

The operator restarts a replica only for *retryable* codes (130, 137, 143 and the user-defined
138; reference pkg/util/train/train_util.go:18-53, controllers read them in
pkg/job_controller/pod.go:282-304).  A Python traceback exits with 1, which is permanent — right
for a bad flag, wrong for what distributed GPU jobs actually die of: a peer went away, so this
rank's collective timed out; the in-kernel watchdog of the fused all-reduce trapped
(`csrc/allreduce.cu`: a spinning wait on a peer's flag becomes `__trap()` after the timeout);
NCCL tore the communicator down; the rendezvous store vanished.  Restarting the pod with the same
rank is the cure for all of those, so they leave with 138.
"""
from __future__ import annotations

import os
import sys
import traceback
from typing import Callable, Optional

from ..utils.train import USER_RETRYABLE_EXIT_CODE

# substrings (lower-case) of exception text that mean "the job is fine, this incarnation is not"
_RETRYABLE_MARKERS = (
    "cuda error", "unspecified launch failure", "device-side assert", "an illegal memory access",
    "cudaerrorlaunchfailure", "cudaerrorecc", "watchdog",
    "nccl", "communicator was aborted", "collective operation timeout",
    "connection reset", "connection refused", "connection closed", "closed by peer", "broken pipe",
    "timeout", "timed out", "gloo/", "processgroup",
    "tcpstore", "the store", "diststoreerror", "distnetworkerror", "distbackenderror", "wait timeout",
    "rendezvous", "peer closed", "unreachable",
    "kdl_err_peer", "peer did not arrive",
)
_PERMANENT_TYPES = (SyntaxError, ImportError, AttributeError, TypeError, KeyError, AssertionError, NotImplementedError)


def classify(exc: BaseException) -> int:
    """Exit code for an exception that escaped a worker's main()."""
    if isinstance(exc, SystemExit):
        code = exc.code
        return code if isinstance(code, int) else (0 if code is None else 1)
    if isinstance(exc, KeyboardInterrupt):
        return 130
    if isinstance(exc, MemoryError):
        return 137                               # what the OOM killer would have reported
    if isinstance(exc, _PERMANENT_TYPES):
        return 1
    if isinstance(exc, (ConnectionError, TimeoutError)):
        return USER_RETRYABLE_EXIT_CODE
    text = f"{type(exc).__name__}: {exc}".lower()
    if "out of memory" in text:
        return 1                                 # the same pod would run out again
    if any(m in text for m in _RETRYABLE_MARKERS):
        return USER_RETRYABLE_EXIT_CODE
    return 1


def run(main: Callable[[], Optional[int]], hard_exit: Optional[bool] = None) -> int:
    """`sys.exit(run(main))`: main's return value, or the classified code of what it raised.
    hard_exit (default: only when running as a worker process, i.e. called from `__main__`):
    on failure, exit the process immediately with that code."""
    if hard_exit is None:
        hard_exit = sys._getframe(1).f_globals.get("__name__") == "__main__"
    try:
        rc = main()
        return int(rc or 0)
    except BaseException as e:                   # noqa: BLE001 - this is the process boundary
        code = classify(e)
        if not isinstance(e, SystemExit):
            traceback.print_exc()
            kind = "retryable" if code in (USER_RETRYABLE_EXIT_CODE, 130, 137, 143) else "permanent"
            print(f"worker failed ({type(e).__name__}); exiting with {code} ({kind} under restartPolicy ExitCode)",
                  file=sys.stderr, flush=True)
            if hard_exit:
                # Leave now.  Normal interpreter shutdown would tear down a process group whose
                # transport threads are mid-failure; gloo then calls std::terminate and the pod
                # reports SIGABRT (134, permanent) instead of the code chosen above.
                try:
                    sys.stdout.flush()
                    sys.stderr.flush()
                finally:
                    os._exit(code)
        return code
