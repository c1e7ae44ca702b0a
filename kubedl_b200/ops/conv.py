"""Implicit-GEMM convolution on the tcgen05 kernel with TMA im2col operand loads (kernel N7).
NHWC bf16 activations, weights [Cout, R, S, Cin]."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import lib as L
from .nn import workspace

L.register_signatures({
    "kdl_conv2d_scratch_bytes": (C.c_size_t, [C.c_int] * 10),
    "kdl_conv2d_bf16": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 9 +
                        [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "kdl_conv2d_fprop_stats": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 9 +
                               [C.c_void_p, C.POINTER(C.c_int), C.c_void_p]),
    "kdl_gemm_max_ctas": (C.c_int, []),
    "kdl_conv2d_dgrad_s2": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 5 + [C.c_void_p]),
    "kdl_conv3x3_halo_supported": (C.c_int, [C.c_int] * 6),
    "kdl_conv3x3_halo_wgrad_scratch_bytes": (C.c_size_t, [C.c_int] * 5),
    "kdl_conv3x3_halo_wgrad_bf16": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 5 +
                                    [C.c_void_p, C.c_size_t, C.c_void_p]),
    "kdl_conv3x3_halo_scratch_bytes": (C.c_size_t, [C.c_int] * 6),
    "kdl_conv3x3_halo_bf16": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 5 +
                              [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]),
})

import os as _os

# 3x3 / stride-1 / pad-1 shapes with C <= 128 go to the halo-resident kernel (csrc/conv3x3_halo.cu);
# KDL_CONV_HALO=0 keeps everything on the TMA-im2col kernel
_HALO = _os.environ.get("KDL_CONV_HALO", "1") != "0"


def halo_supported(mode, N, H, W, Cin, Cout) -> bool:
    return bool(L.lib().kdl_conv3x3_halo_supported(mode, N, H, W, Cin, Cout))


def _run_halo(mode, in0, w, out, N, H, W, Cin, Cout, addend=None, relu=False):
    lib = L.lib()
    nb = lib.kdl_conv3x3_halo_scratch_bytes(mode, N, H, W, Cin, Cout)
    # the flipped filter copy is consumed by the very next launch on the same stream: one buffer serves all layers
    ws = workspace(nb, out.device, key="conv_wt") if nb else None
    L.check(lib.kdl_conv3x3_halo_bf16(mode, in0.data_ptr(), w.data_ptr(), out.data_ptr(), N, H, W, Cin, Cout,
                                      L.ptr(addend), int(relu), L.ptr(ws), ws.numel() if ws is not None else 0,
                                      L.stream_ptr()), "kdl_conv3x3_halo_bf16")
    L.count(2 if nb else 1)
    return out


def _chk(t):
    assert t.dtype == torch.bfloat16 and t.is_cuda and t.is_contiguous(), (t.dtype, t.shape)


def _run(mode, in0, in1, out, N, H, W, Cin, Cout, R, S, stride, pad, addend=None):
    lib = L.lib()
    nb = lib.kdl_conv2d_scratch_bytes(mode, N, H, W, Cin, Cout, R, S, stride, pad)
    ws = workspace(nb, out.device, key="gemm_splitk") if nb else None
    L.check(lib.kdl_conv2d_bf16(mode, in0.data_ptr(), in1.data_ptr(), out.data_ptr(), N, H, W, Cin, Cout, R, S,
                                stride, pad, L.ptr(addend), L.ptr(ws), ws.numel() if ws is not None else 0,
                                L.stream_ptr()), "kdl_conv2d_bf16")
    L.count(2 if (mode == 2 and nb) else 1)
    return out


def fprop(x: torch.Tensor, w: torch.Tensor, stride: int, pad: int, halo: bool = True, want_stats: bool = False):
    """y = conv(x, w).  want_stats: returns (y, stats) where stats = (partial, nparts) for ops.nn.bn_fwd(stats=...) when
    the convolution ran on the TMA-im2col kernel (BatchNorm column statistics from its epilogue), else (y, None)."""
    _chk(x); _chk(w)
    N, H, W, Cin = x.shape
    Cout, R, S, _ = w.shape
    OH, OW = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    y = torch.empty(N, OH, OW, Cout, dtype=torch.bfloat16, device=x.device)
    if _HALO and halo and R == 3 and S == 3 and stride == 1 and pad == 1 and halo_supported(0, N, H, W, Cin, Cout):
        y = _run_halo(0, x, w, y, N, H, W, Cin, Cout)
        return (y, None) if want_stats else y
    if want_stats and Cout <= 2048:
        lib = L.lib()
        partial = torch.empty(lib.kdl_gemm_max_ctas(), 2, Cout, dtype=torch.float32, device=x.device)
        nparts = C.c_int(0)
        L.check(lib.kdl_conv2d_fprop_stats(x.data_ptr(), w.data_ptr(), y.data_ptr(), N, H, W, Cin, Cout, R, S, stride, pad,
                                           partial.data_ptr(), C.byref(nparts), L.stream_ptr()), "kdl_conv2d_fprop_stats")
        L.count(1)
        return y, (partial, nparts.value)
    y = _run(0, x, w, y, N, H, W, Cin, Cout, R, S, stride, pad)
    return (y, None) if want_stats else y


def dgrad(dy: torch.Tensor, w: torch.Tensor, in_hw, stride: int, pad: int,
          addend: Optional[torch.Tensor] = None, halo: bool = True) -> torch.Tensor:
    _chk(dy); _chk(w)
    N, OH, OW, Cout = dy.shape
    _, R, S, Cin = w.shape
    H, W = in_hw
    dx = torch.empty(N, H, W, Cin, dtype=torch.bfloat16, device=dy.device)
    if _HALO and halo and R == 3 and S == 3 and stride == 1 and pad == 1 and halo_supported(1, N, H, W, Cin, Cout):
        return _run_halo(1, dy, w, dx, N, H, W, Cin, Cout, addend=addend)
    return _run(1, dy, w, dx, N, H, W, Cin, Cout, R, S, stride, pad, addend)


def dgrad_s2_supported(in_hw, Cin: int, Cout: int, R: int, S: int, stride: int, pad: int) -> bool:
    H, W = in_hw
    return R == 3 and S == 3 and stride == 2 and pad == 1 and H % 2 == 0 and W % 2 == 0 and Cin % 64 == 0 and Cout % 64 == 0


def dgrad_s2(dy: torch.Tensor, w: torch.Tensor, in_hw) -> torch.Tensor:
    """Input gradient of a 3x3 / stride-2 / pad-1 convolution: four parity sub-convolutions, one persistent
    launch of the im2col kernel."""
    _chk(dy); _chk(w)
    N, OH, OW, Cout = dy.shape
    _, R, S, Cin = w.shape
    H, W = in_hw
    assert dgrad_s2_supported(in_hw, Cin, Cout, R, S, 2, 1) and OH == H // 2 and OW == W // 2
    dx = torch.empty(N, H, W, Cin, dtype=torch.bfloat16, device=dy.device)
    L.check(L.lib().kdl_conv2d_dgrad_s2(dy.data_ptr(), w.data_ptr(), dx.data_ptr(), N, H, W, Cin, Cout, L.stream_ptr()),
            "kdl_conv2d_dgrad_s2")
    L.count(1)
    return dx


def wgrad(dy: torch.Tensor, x: torch.Tensor, R: int, S: int, stride: int, pad: int,
          out: Optional[torch.Tensor] = None, halo: bool = True) -> torch.Tensor:
    _chk(dy); _chk(x)
    N, H, W, Cin = x.shape
    Cout = dy.shape[-1]
    dw = out if out is not None else torch.empty(Cout, R, S, Cin, dtype=torch.bfloat16, device=x.device)
    if _HALO and halo and R == 3 and S == 3 and stride == 1 and pad == 1:
        lib = L.lib()
        nb = lib.kdl_conv3x3_halo_wgrad_scratch_bytes(N, H, W, Cin, Cout)
        if nb:
            ws = workspace(nb, dw.device, key="gemm_splitk")
            L.check(lib.kdl_conv3x3_halo_wgrad_bf16(dy.data_ptr(), x.data_ptr(), dw.data_ptr(), N, H, W, Cin, Cout,
                                                    ws.data_ptr(), ws.numel(), L.stream_ptr()),
                    "kdl_conv3x3_halo_wgrad_bf16")
            L.count(2)
            return dw
    return _run(2, dy, x, dw, N, H, W, Cin, Cout, R, S, stride, pad)
