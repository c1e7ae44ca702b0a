"""ctypes binding of libkubedl_b200.so (C ABI in include/kubedl_b200.h).

The library is the only compute path of the worker runtime: there is no eager fallback.  If
the .so cannot be loaded the first kernel call raises, so a GPU test can never silently pass
on a PyTorch re-implementation.  `launch_count()` reports how many of OUR kernels have been
launched (bench.py's `gpu_launches`).
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Optional

from . import build as _build

_lock = threading.Lock()
_lib: Optional[C.CDLL] = None
_launches = 0

c_void_p, c_int, c_int64, c_float, c_size_t, c_uint, c_ull = (
    C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t, C.c_uint, C.c_ulonglong)

_SIGS = {
    "kdl_last_error": (C.c_char_p, []),
    "kdl_version": (c_int, []),
    "kdl_bn_workspace_bytes": (c_size_t, [c_int64, c_int]),
    "kdl_bn_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p,
                           c_void_p, c_void_p, c_float, c_float, c_int, c_void_p, c_void_p,
                           c_size_t, c_void_p]),
    "kdl_bn_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int,
                           c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_size_t,
                           c_void_p]),
    "kdl_bn_fwd_mask": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p,
                                c_void_p, c_void_p, c_float, c_float, c_int, c_void_p, c_void_p, c_void_p,
                                c_size_t, c_void_p]),
    "kdl_bn_bwd_mask": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int,
                                c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "kdl_bn_fwd_from_stats_mask": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p,
                                           c_void_p, c_void_p, c_float, c_float, c_int, c_void_p, c_void_p, c_int,
                                           c_void_p, c_void_p, c_size_t, c_void_p]),
    "kdl_bn_fwd_from_stats": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_float, c_float, c_int, c_void_p, c_void_p, c_int,
                                      c_void_p, c_size_t, c_void_p]),
    "kdl_u8_normalize": (c_int, [c_void_p, c_void_p, c_int64, c_int, C.POINTER(c_float),
                                 C.POINTER(c_float), c_void_p]),
    "kdl_stem_im2col_u8": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, C.POINTER(c_float),
                                   C.POINTER(c_float), c_void_p]),
    "kdl_bn_bwd_pooled": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "kdl_bn_finalize_from_stats": (c_int, [c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float,
                                           c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "kdl_maxpool3x3s2_affine_relu_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                                 c_void_p]),
    "kdl_stem_conv_supported": (c_int, [c_int, c_int]),
    "kdl_stem_conv_max_parts": (c_int, []),
    "kdl_stem_conv_wgrad_scratch_bytes": (c_size_t, []),
    "kdl_stem_conv_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, C.POINTER(c_float),
                                  C.POINTER(c_float), c_void_p, C.POINTER(c_int), c_void_p]),
    "kdl_stem_conv_wgrad": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, C.POINTER(c_float),
                                    C.POINTER(c_float), c_void_p, c_size_t, c_void_p]),
    "kdl_stem_s2d_image_bytes": (c_size_t, [c_int]),
    "kdl_stem_s2d_wgrad_scratch_bytes": (c_size_t, []),
    "kdl_stem_s2d_pack": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, C.POINTER(c_float), C.POINTER(c_float),
                                  c_void_p]),
    "kdl_stem_s2d_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p,
                                 C.POINTER(c_int), c_void_p]),
    "kdl_stem_s2d_wgrad": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_size_t, c_void_p]),
    "kdl_softmax_xent_workspace_bytes": (c_size_t, [c_int]),
    "kdl_softmax_xent": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                 c_int, c_float, c_void_p, c_size_t, c_void_p]),
    "kdl_sgd_momentum": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64,
                                 c_float, c_float, c_float, c_float, c_void_p]),
    "kdl_add_bf16": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "kdl_maxpool3x3s2_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                     c_void_p]),
    "kdl_maxpool3x3s2_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                     c_void_p]),
    "kdl_avgpool_fwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "kdl_avgpool_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "kdl_allreduce_flags_bytes": (c_size_t, [c_int]),
    "kdl_allreduce_fused": (c_int, [C.POINTER(c_void_p), C.POINTER(c_void_p), C.POINTER(c_void_p),
                                    c_void_p, c_void_p, c_int, c_int, c_int64, c_int, c_int,
                                    c_float, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                    c_ull, c_void_p]),
    "kdl_symm_barrier": (c_int, [C.POINTER(c_void_p), c_int, c_int, c_void_p, c_int, c_int, c_int, c_void_p,
                                 c_ull, c_void_p]),
    "kdl_symm_malloc": (c_int, [c_size_t, C.POINTER(c_void_p)]),
    "kdl_symm_free": (c_int, [c_void_p]),
    "kdl_ipc_get_handle": (c_int, [c_void_p, c_void_p]),
    "kdl_ipc_open_handle": (c_int, [c_void_p, C.POINTER(c_void_p)]),
    "kdl_ipc_close_handle": (c_int, [c_void_p]),
    "kdl_enable_peer_access": (c_int, [c_int]),
    "kdl_vmm_supported": (c_int, [c_int, C.POINTER(c_int), C.POINTER(c_int)]),
    "kdl_vmm_granularity": (c_size_t, [c_int, c_int]),
    "kdl_vmm_alloc": (c_int, [c_int, c_size_t, C.POINTER(c_ull), C.POINTER(c_void_p),
                              C.POINTER(c_int)]),
    "kdl_vmm_import": (c_int, [c_int, c_int, c_size_t, C.POINTER(c_ull), C.POINTER(c_void_p)]),
    "kdl_vmm_release": (c_int, [c_ull, c_void_p, c_size_t]),
    "kdl_mc_create": (c_int, [c_int, c_size_t, C.POINTER(c_ull), C.POINTER(c_int)]),
    "kdl_mc_import": (c_int, [c_int, C.POINTER(c_ull)]),
    "kdl_mc_add_device": (c_int, [c_ull, c_int]),
    "kdl_mc_bind_and_map": (c_int, [c_ull, c_ull, c_int, c_size_t, C.POINTER(c_void_p)]),
    "kdl_mc_release": (c_int, [c_ull, c_void_p, c_size_t]),
}


class KernelError(RuntimeError):
    pass


def register_signatures(sigs: dict) -> None:
    """Let other op modules (gemm, hist, embedding) add their entry points."""
    _SIGS.update(sigs)
    global _lib
    if _lib is not None:
        _apply(_lib, sigs)


def _apply(lib: C.CDLL, sigs: dict) -> None:
    for name, (res, args) in sigs.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:  # stale .so
            raise KernelError(f"libkubedl_b200.so lacks {name}; rebuild (python -m kubedl_b200.ops.build)") from e
        fn.restype = res
        fn.argtypes = args


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                path = _build.ensure_built()
                l = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
                _apply(l, _SIGS)
                _lib = l
    return _lib


def lib_path() -> str:
    return str(_build.LIB_PATH)


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().kdl_last_error()
        raise KernelError(f"{what} failed (rc={rc}): {msg.decode() if msg else '?'}")


def count(n: int = 1) -> None:
    global _launches
    _launches += n


def launch_count() -> int:
    return _launches


def reset_launch_count() -> None:
    global _launches
    _launches = 0


def stream_ptr() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


def ptr(t) -> Optional[int]:
    return None if t is None else t.data_ptr()
