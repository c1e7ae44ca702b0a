"""tcgen05 GEMM front-end (csrc/gemm_tcgen05.cu, kernel N6).

    g = Gemm(device)
    y  = g.nt(x, w)            # y[M,N]  = x[M,K]  @ w[N,K]^T      (forward of 1x1 conv / linear)
    dx = g.nn(dy, w)           # dx[M,K'] = dy[M,N'] @ w[N',K']     (dgrad)
    dw = g.tn(dy, x, out=gw)   # dw[N',K'] = dy[M,N']^T @ x[M,K']   (wgrad, split-K over M)

All operands are row-major bf16 tensors; the transposes are expressed through UMMA
shared-memory descriptors (MN-major operands), never by copying.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import lib as L
from .nn import workspace

L.register_signatures({
    "kdl_gemm_splits": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "kdl_gemm_scratch_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "kdl_gemm_max_ctas": (C.c_int, []),
    "kdl_gemm_bf16_ex": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                   C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                   C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.POINTER(C.c_int),
                                   C.c_void_p]),
    "kdl_gemm_bf16_subadd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "kdl_gemm_bf16": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]),
    "kdl_gemm_bf16_maskadd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                        C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
})


def _chk(t: torch.Tensor):
    assert t.dtype == torch.bfloat16 and t.is_cuda and t.dim() == 2 and t.stride(1) == 1, \
        f"need row-major CUDA bf16 matrix, got {t.dtype} {tuple(t.shape)} strides {t.stride()}"


class Gemm:
    def __init__(self, device=None):
        self.lib = L.lib()
        self.device = device

    def nt_stats(self, a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None):
        """y = a @ b^T plus, from the same epilogue, per-CTA column sum / sum-of-squares of y (as
        stored, i.e. bf16-rounded).  Returns (y, (partial[nparts_max, 2, N] fp32, nparts)) for
        ops.nn.bn_fwd(stats=...)."""
        _chk(a); _chk(b)
        M, K = a.shape
        N = b.shape[0]
        if out is None:
            out = torch.empty(M, N, dtype=torch.bfloat16, device=a.device)
        partial = torch.empty(self.lib.kdl_gemm_max_ctas(), 2, N, dtype=torch.float32, device=a.device)
        nparts = C.c_int(0)
        rc = self.lib.kdl_gemm_bf16_ex(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, 0, 0, a.stride(0),
                                       b.stride(0), out.stride(0), None, None, 0, 1, None, 0, partial.data_ptr(),
                                       C.byref(nparts), L.stream_ptr())
        L.check(rc, "kdl_gemm_bf16_ex")
        L.count(1)
        return out, (partial, nparts.value)

    def _run(self, A, B, out, M, N, K, a_mn, b_mn, bias, addend, relu, split_k):
        lib = self.lib
        splits = lib.kdl_gemm_splits(M, N, K, split_k)
        ws = None
        nbytes = 0
        if splits > 1:
            nbytes = lib.kdl_gemm_scratch_bytes(M, N, K, split_k)
            ws = workspace(nbytes, out.device, key="gemm_splitk")
        rc = lib.kdl_gemm_bf16(A.data_ptr(), B.data_ptr(), out.data_ptr(), M, N, K, a_mn, b_mn,
                               A.stride(0), B.stride(0), out.stride(0), L.ptr(bias), L.ptr(addend),
                               int(relu), split_k, L.ptr(ws), ws.numel() if ws is not None else 0,
                               L.stream_ptr())
        L.check(rc, "kdl_gemm_bf16")
        L.count(2 if splits > 1 else 1)
        return out

    def nt(self, a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None,
           bias: Optional[torch.Tensor] = None, addend: Optional[torch.Tensor] = None,
           relu: bool = False) -> torch.Tensor:
        _chk(a); _chk(b)
        M, K = a.shape
        N, K2 = b.shape
        assert K == K2
        if out is None:
            out = torch.empty(M, N, dtype=torch.bfloat16, device=a.device)
        return self._run(a, b, out, M, N, K, 0, 0, bias, addend, relu, 1)

    def nn(self, a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None,
           addend: Optional[torch.Tensor] = None, addend_hw=None,
           addend_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        """out = a @ b (+ addend).  addend_hw=(H, W): the rows of `out` are the pixels of an [n, H, W] map
        and `addend` is the compact [n*H/2*W/2, N] gradient of a 1x1 / stride-2 branch, added at even
        (h, w) only — the up-sampled (three quarters zero) tensor is never materialised.
        addend_mask: uint8 [M*N/8] ReLU bit mask (ops.nn.bn_fwd(relu_mask=...)); the addend is added only
        where its bit is set, i.e. out = a @ b + relu'(.) * addend without a masked copy of the addend."""
        _chk(a); _chk(b)
        M, K = a.shape
        K2, N = b.shape
        assert K == K2
        if out is None:
            out = torch.empty(M, N, dtype=torch.bfloat16, device=a.device)
        if addend_mask is not None:
            _chk(addend)
            assert addend_hw is None and addend.shape == (M, N) and addend.stride(0) == N and out.stride(0) == N
            assert addend_mask.dtype == torch.uint8 and addend_mask.numel() * 8 == M * N
            L.check(self.lib.kdl_gemm_bf16_maskadd(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, 0, 1, a.stride(0),
                                                   b.stride(0), N, addend.data_ptr(), addend_mask.data_ptr(),
                                                   L.stream_ptr()), "kdl_gemm_bf16_maskadd")
            L.count(1)
            return out
        if addend_hw is not None:
            H, W = addend_hw
            _chk(addend)
            assert addend.shape == (M // 4, N) and addend.stride(0) == out.stride(0)
            L.check(self.lib.kdl_gemm_bf16_subadd(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, 0, 1, a.stride(0),
                                                  b.stride(0), out.stride(0), addend.data_ptr(), H, W, L.stream_ptr()),
                    "kdl_gemm_bf16_subadd")
            L.count(1)
            return out
        return self._run(a, b, out, M, N, K, 0, 1, None, addend, False, 1)

    def tn(self, a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None,
           split_k: int = 0) -> torch.Tensor:
        """out[Ma, Nb] = a[K, Ma]^T @ b[K, Nb]."""
        _chk(a); _chk(b)
        K, M = a.shape
        K2, N = b.shape
        assert K == K2
        if out is None:
            out = torch.empty(M, N, dtype=torch.bfloat16, device=a.device)
        return self._run(a, b, out, M, N, K, 1, 1, None, None, False, split_k)
