"""Tensor-level wrappers over the fused memory-bound kernels (csrc/batchnorm.cu,
csrc/elementwise.cu).  All activations are NHWC bf16; a [N,H,W,C] tensor is passed as the
[M, C] matrix it is in memory.  Each wrapper launches on torch's current stream, so the calls
can be captured into CUDA graphs and ordered against cuDNN/cuBLAS work issued through torch.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import lib as L

_workspaces: dict = {}


def workspace(nbytes: int, device=None, key: str = "default") -> torch.Tensor:
    """Zero-initialised scratch arena, grown on demand, one per (device, key)."""
    device = torch.device(device if device is not None else torch.cuda.current_device())
    if device.type == "cuda" and device.index is None:
        device = torch.device("cuda", torch.cuda.current_device())
    k = (device, key)
    ws = _workspaces.get(k)
    if ws is None or ws.numel() < nbytes:
        ws = torch.zeros(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[k] = ws
    return ws


def _chk_bf16(*ts):
    for t in ts:
        if t is not None:
            assert t.dtype == torch.bfloat16 and t.is_contiguous() and t.is_cuda, \
                f"need contiguous CUDA bf16, got {t.dtype} {tuple(t.shape)}"


def bn_fwd(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor,
           running_mean: Optional[torch.Tensor], running_var: Optional[torch.Tensor],
           res: Optional[torch.Tensor] = None, relu: bool = True, eps: float = 1e-5,
           momentum: float = 0.1, out: Optional[torch.Tensor] = None,
           save: Optional[torch.Tensor] = None, stats=None,
           relu_mask: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """y = relu(bn_train(x) + res) on x:[M,C] (or [N,H,W,C]) bf16. Returns (y, save_mean_rstd[2,C]).
    `stats` = (partial, nparts) from Gemm.nt_stats: the statistics pass over x is skipped.
    `relu_mask` (uint8 [M*C/8]) receives one bit per element, "positive before the ReLU", for
    bn_bwd(relu_mask=...) — 1/16 of the bytes of y."""
    _chk_bf16(x, res)
    Cn = x.shape[-1]
    M = x.numel() // Cn
    y = out if out is not None else torch.empty_like(x)
    if save is None:
        save = torch.empty(2, Cn, dtype=torch.float32, device=x.device)
    lib = L.lib()
    need = lib.kdl_bn_workspace_bytes(M, Cn)
    if need == 0:
        raise L.KernelError(f"bn_fwd: unsupported shape M={M} C={Cn}")
    ws = workspace(need, x.device)
    if relu_mask is not None:
        assert relu_mask.dtype == torch.uint8 and relu_mask.numel() * 8 == x.numel() and relu_mask.is_contiguous()
    if stats is not None:
        partial, nparts = stats
        L.check(lib.kdl_bn_fwd_from_stats_mask(x.data_ptr(), L.ptr(res), y.data_ptr(), M, Cn, gamma.data_ptr(),
                                               beta.data_ptr(), L.ptr(running_mean), L.ptr(running_var), eps,
                                               momentum, int(relu), save.data_ptr(), partial.data_ptr(), nparts,
                                               L.ptr(relu_mask), ws.data_ptr(), ws.numel(), L.stream_ptr()),
                "kdl_bn_fwd_from_stats_mask")
        L.count(2)
        return y, save
    L.check(lib.kdl_bn_fwd_mask(x.data_ptr(), L.ptr(res), y.data_ptr(), M, Cn, gamma.data_ptr(),
                                beta.data_ptr(), L.ptr(running_mean), L.ptr(running_var), eps, momentum,
                                int(relu), save.data_ptr(), L.ptr(relu_mask), ws.data_ptr(), ws.numel(),
                                L.stream_ptr()), "kdl_bn_fwd_mask")
    L.count(2)
    return y, save


def bn_bwd(dy: torch.Tensor, x: torch.Tensor, y: Optional[torch.Tensor], gamma: torch.Tensor,
           save: torch.Tensor, relu: bool, dgamma: torch.Tensor, dbeta: torch.Tensor,
           dx: Optional[torch.Tensor] = None, dres: Optional[torch.Tensor] = None,
           beta: Optional[torch.Tensor] = None, relu_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Backward of bn_fwd.  dgamma/dbeta: bf16 [C] (views into the gradient arena).  If `dres`
    is given it receives dz = dy * relu_mask (the gradient of the residual input); it may alias
    dy.  With relu and y=None (layers without a residual input) the mask is recomputed from x with
    the forward affine (`beta` required) and y is never read; with `relu_mask` (the bit mask bn_fwd
    saved) the mask bits are read instead of y."""
    _chk_bf16(dy, x, y, dx, dres, dgamma, dbeta)
    Cn = x.shape[-1]
    M = x.numel() // Cn
    if dx is None:
        dx = torch.empty_like(x)
    lib = L.lib()
    need = lib.kdl_bn_workspace_bytes(M, Cn)
    ws = workspace(need, x.device)
    if relu and relu_mask is not None:
        L.check(lib.kdl_bn_bwd_mask(dy.data_ptr(), x.data_ptr(), relu_mask.data_ptr(), dx.data_ptr(), L.ptr(dres), M, Cn,
                                    gamma.data_ptr(), save.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(),
                                    ws.data_ptr(), ws.numel(), L.stream_ptr()), "kdl_bn_bwd_mask")
        L.count(2)
        return dx
    L.check(lib.kdl_bn_bwd(dy.data_ptr(), x.data_ptr(), L.ptr(y), dx.data_ptr(), L.ptr(dres), M, Cn,
                           gamma.data_ptr(), L.ptr(beta), save.data_ptr(), int(relu), dgamma.data_ptr(),
                           dbeta.data_ptr(), ws.data_ptr(), ws.numel(), L.stream_ptr()),
            "kdl_bn_bwd")
    L.count(2)
    return dx


_IMAGENET_MEAN = (C.c_float * 3)(0.485, 0.456, 0.406)
_IMAGENET_STD = (C.c_float * 3)(0.229, 0.224, 0.225)


def u8_normalize(img_u8: torch.Tensor, cpad: int = 3, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """uint8 NHWC [N,H,W,3] -> normalised bf16 NHWC [N,H,W,cpad]."""
    assert img_u8.dtype == torch.uint8 and img_u8.shape[-1] == 3 and img_u8.is_contiguous()
    npix = img_u8.numel() // 3
    if out is None:
        out = torch.empty(*img_u8.shape[:-1], cpad, dtype=torch.bfloat16, device=img_u8.device)
    L.check(L.lib().kdl_u8_normalize(img_u8.data_ptr(), out.data_ptr(), npix, cpad, _IMAGENET_MEAN,
                                     _IMAGENET_STD, L.stream_ptr()), "kdl_u8_normalize")
    L.count(1)
    return out


STEM_KW = 24            # a filter row (7 pixels x 3 channels = 21 values) padded to 24
STEM_KP = 7 * STEM_KW   # 168 patch columns


def stem_weight_padded(w: torch.Tensor) -> torch.Tensor:
    """[64, 7, 7, 3] stem filter -> [64, 168] in the patch-column layout of stem_im2col."""
    wp = torch.zeros(w.shape[0], 7, STEM_KW, dtype=w.dtype, device=w.device)
    wp[:, :, :21] = w.reshape(w.shape[0], 7, 21)
    return wp.view(w.shape[0], STEM_KP)


def stem_weight_unpad(gw: torch.Tensor) -> torch.Tensor:
    """[64, 168] gradient in patch-column layout -> [64, 147]."""
    return gw.view(gw.shape[0], 7, STEM_KW)[:, :, :21].reshape(gw.shape[0], 147)


def stem_im2col(img_u8: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """uint8 NHWC [N,H,W,3] -> normalised bf16 patches [N*OH*OW, 168] of the 7x7/s2/p3 stem conv
    (7 filter rows x 24 columns, 21 used)."""
    assert img_u8.dtype == torch.uint8 and img_u8.shape[-1] == 3 and img_u8.is_contiguous()
    N, H, W, _ = img_u8.shape
    OH, OW = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    if out is None:
        out = torch.empty(N * OH * OW, STEM_KP, dtype=torch.bfloat16, device=img_u8.device)
    L.check(L.lib().kdl_stem_im2col_u8(img_u8.data_ptr(), out.data_ptr(), N, H, W, STEM_KP, _IMAGENET_MEAN,
                                       _IMAGENET_STD, L.stream_ptr()), "kdl_stem_im2col_u8")
    L.count(1)
    return out


def stem_conv_supported(H: int, W: int) -> bool:
    return bool(L.lib().kdl_stem_conv_supported(H, W))


def stem_conv_fwd(img_u8: torch.Tensor, w_padded: torch.Tensor, want_stats: bool = False):
    """7x7/s2/p3 stem convolution straight from the uint8 NHWC image (csrc/stem_conv.cu): no patch matrix
    in memory.  w_padded: [64, 168] (stem_weight_padded).  Returns (y [N,112,112,64], stats or None) with
    stats = (partial, nparts) for bn_fwd(stats=...)."""
    assert img_u8.dtype == torch.uint8 and img_u8.shape[-1] == 3 and img_u8.is_contiguous()
    assert w_padded.shape == (64, STEM_KP) and w_padded.is_contiguous()
    N, H, W, _ = img_u8.shape
    lib = L.lib()
    y = torch.empty(N, H // 2, W // 2, 64, dtype=torch.bfloat16, device=img_u8.device)
    partial, nparts = None, C.c_int(0)
    if want_stats:
        partial = torch.empty(lib.kdl_stem_conv_max_parts(), 2, 64, dtype=torch.float32, device=img_u8.device)
    L.check(lib.kdl_stem_conv_fwd(img_u8.data_ptr(), w_padded.data_ptr(), y.data_ptr(), N, H, W, _IMAGENET_MEAN,
                                  _IMAGENET_STD, L.ptr(partial), C.byref(nparts), L.stream_ptr()), "kdl_stem_conv_fwd")
    L.count(1)
    return y, ((partial, nparts.value) if want_stats else None)


def stem_conv_wgrad(img_u8: torch.Tensor, dy: torch.Tensor) -> torch.Tensor:
    """Weight gradient of the stem, [64, 168] in the padded column layout (stem_weight_unpad)."""
    N, H, W, _ = img_u8.shape
    assert dy.dtype == torch.bfloat16 and dy.is_contiguous() and dy.numel() == N * (H // 2) * (W // 2) * 64
    lib = L.lib()
    ws = workspace(lib.kdl_stem_conv_wgrad_scratch_bytes(), img_u8.device, key="stem_wgrad")
    gw = torch.empty(64, STEM_KP, dtype=torch.bfloat16, device=img_u8.device)
    L.check(lib.kdl_stem_conv_wgrad(img_u8.data_ptr(), dy.data_ptr(), gw.data_ptr(), N, H, W, _IMAGENET_MEAN,
                                    _IMAGENET_STD, ws.data_ptr(), ws.numel(), L.stream_ptr()), "kdl_stem_conv_wgrad")
    L.count(2)
    return gw


def stem_s2d_pack(img_u8: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Normalised, zero-bordered 2x2 space-to-depth form of the uint8 NHWC image: [N, 115, 115, 16] bf16
    (csrc/stem_s2d.cu).  Written once per step; both stem kernels read their operand tiles from it with TMA."""
    assert img_u8.dtype == torch.uint8 and img_u8.shape[-1] == 3 and img_u8.is_contiguous()
    N, H, W, _ = img_u8.shape
    if out is None:
        out = torch.empty(N, 115, 115, 16, dtype=torch.bfloat16, device=img_u8.device)
    assert out.numel() * 2 == L.lib().kdl_stem_s2d_image_bytes(N)
    L.check(L.lib().kdl_stem_s2d_pack(img_u8.data_ptr(), out.data_ptr(), N, H, W, _IMAGENET_MEAN, _IMAGENET_STD,
                                      L.stream_ptr()), "kdl_stem_s2d_pack")
    L.count(1)
    return out


def stem_s2d_fwd(s2d: torch.Tensor, w_padded: torch.Tensor, want_stats: bool = False):
    """The stem convolution as a 4x4 convolution on the space-to-depth image.  Same contract as stem_conv_fwd."""
    assert s2d.dtype == torch.bfloat16 and s2d.shape[1:] == (115, 115, 16) and s2d.is_contiguous()
    assert w_padded.shape == (64, STEM_KP) and w_padded.is_contiguous()
    N = s2d.shape[0]
    lib = L.lib()
    y = torch.empty(N, 112, 112, 64, dtype=torch.bfloat16, device=s2d.device)
    w2 = workspace(64 * 256 * 2, s2d.device, key="stem_s2d_w2")
    partial, nparts = None, C.c_int(0)
    if want_stats:
        partial = torch.empty(lib.kdl_stem_conv_max_parts(), 2, 64, dtype=torch.float32, device=s2d.device)
    L.check(lib.kdl_stem_s2d_fwd(s2d.data_ptr(), w_padded.data_ptr(), w2.data_ptr(), y.data_ptr(), N, 224, 224,
                                 L.ptr(partial), C.byref(nparts), L.stream_ptr()), "kdl_stem_s2d_fwd")
    L.count(2)
    return y, ((partial, nparts.value) if want_stats else None)


def stem_s2d_wgrad(s2d: torch.Tensor, dy: torch.Tensor) -> torch.Tensor:
    """Weight gradient of the stem from the space-to-depth image, [64, 168] in the padded column layout."""
    N = s2d.shape[0]
    assert dy.dtype == torch.bfloat16 and dy.is_contiguous() and dy.numel() == N * 112 * 112 * 64
    lib = L.lib()
    ws = workspace(lib.kdl_stem_s2d_wgrad_scratch_bytes(), s2d.device, key="stem_s2d_wgrad")
    gw = torch.empty(64, STEM_KP, dtype=torch.bfloat16, device=s2d.device)
    L.check(lib.kdl_stem_s2d_wgrad(s2d.data_ptr(), dy.data_ptr(), gw.data_ptr(), N, 224, 224, ws.data_ptr(), ws.numel(),
                                   L.stream_ptr()), "kdl_stem_s2d_wgrad")
    L.count(2)
    return gw


def softmax_xent(logits: torch.Tensor, labels: torch.Tensor, V: Optional[int] = None,
                 grad_scale: Optional[float] = None, dlogits: Optional[torch.Tensor] = None,
                 loss: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Mean softmax cross-entropy over rows of logits[B, ldv] (first V columns valid) and its
    gradient (bf16, already scaled by grad_scale, default 1/B)."""
    assert logits.dim() == 2 and logits.is_contiguous() and labels.dtype == torch.int64
    B, ldv = logits.shape
    V = V if V is not None else ldv
    if grad_scale is None:
        grad_scale = 1.0 / B
    if dlogits is None:
        dlogits = torch.empty(B, ldv, dtype=torch.bfloat16, device=logits.device)
    if loss is None:
        loss = torch.empty(1, dtype=torch.float32, device=logits.device)
    lib = L.lib()
    need = lib.kdl_softmax_xent_workspace_bytes(B)
    ws = workspace(need, logits.device, key="xent")
    L.check(lib.kdl_softmax_xent(logits.data_ptr(), int(logits.dtype == torch.float32),
                                 labels.data_ptr(), dlogits.data_ptr(), loss.data_ptr(), B, V, ldv,
                                 grad_scale, ws.data_ptr(), ws.numel(), L.stream_ptr()),
            "kdl_softmax_xent")
    L.count(1)
    return loss, dlogits


def sgd_momentum(p32: torch.Tensor, mom: torch.Tensor, g16: torch.Tensor, p16: torch.Tensor,
                 lr: float, momentum: float = 0.9, weight_decay: float = 0.0,
                 wd_count: Optional[int] = None, grad_scale: float = 1.0) -> None:
    n = p32.numel()
    assert mom.numel() == n and g16.numel() == n and p16.numel() == n
    assert p32.dtype == torch.float32 and g16.dtype == torch.bfloat16 and p16.dtype == torch.bfloat16
    L.check(L.lib().kdl_sgd_momentum(p32.data_ptr(), mom.data_ptr(), g16.data_ptr(), p16.data_ptr(),
                                     n, n if wd_count is None else wd_count, lr, momentum,
                                     weight_decay, grad_scale, L.stream_ptr()), "kdl_sgd_momentum")
    L.count(1)


def add_(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    _chk_bf16(a, b, out)
    out = a if out is None else out
    L.check(L.lib().kdl_add_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), a.numel(),
                                 L.stream_ptr()), "kdl_add_bf16")
    L.count(1)
    return out


def maxpool3x3s2_fwd(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    _chk_bf16(x)
    N, H, W, Cn = x.shape
    OH, OW = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    y = torch.empty(N, OH, OW, Cn, dtype=torch.bfloat16, device=x.device)
    idx = torch.empty(N, OH, OW, Cn, dtype=torch.uint8, device=x.device)
    L.check(L.lib().kdl_maxpool3x3s2_fwd(x.data_ptr(), y.data_ptr(), idx.data_ptr(), N, H, W, Cn,
                                         L.stream_ptr()), "kdl_maxpool3x3s2_fwd")
    L.count(1)
    return y, idx


def bn_relu_maxpool_fwd(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, running_mean, running_var,
                        save: torch.Tensor, stats, eps: float = 1e-5, momentum: float = 0.1):
    """maxpool3x3s2(relu(bn_train(x))) in one pass over x [N,H,W,C] (a raw convolution output whose column
    statistics `stats` = (partial, nparts) came from the producing kernel's epilogue).  The normalised
    activation is never materialised.  Returns (pooled, idx); `save` receives mean / rstd."""
    _chk_bf16(x)
    N, H, W, Cn = x.shape
    partial, nparts = stats
    lib = L.lib()
    coef = torch.empty(2, Cn, dtype=torch.float32, device=x.device)
    L.check(lib.kdl_bn_finalize_from_stats(N * H * W, Cn, gamma.data_ptr(), beta.data_ptr(), L.ptr(running_mean),
                                           L.ptr(running_var), eps, momentum, save.data_ptr(), partial.data_ptr(),
                                           nparts, coef.data_ptr(), L.stream_ptr()), "kdl_bn_finalize_from_stats")
    OH, OW = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    y = torch.empty(N, OH, OW, Cn, dtype=torch.bfloat16, device=x.device)
    idx = torch.empty(N, OH, OW, Cn, dtype=torch.uint8, device=x.device)
    L.check(lib.kdl_maxpool3x3s2_affine_relu_fwd(x.data_ptr(), coef.data_ptr(), y.data_ptr(), idx.data_ptr(), N, H, W,
                                                 Cn, L.stream_ptr()), "kdl_maxpool3x3s2_affine_relu_fwd")
    L.count(2)
    return y, idx


def bn_bwd_pooled(dp: torch.Tensor, idx: torch.Tensor, x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor,
                  save: torch.Tensor, dgamma: torch.Tensor, dbeta: torch.Tensor) -> torch.Tensor:
    """Backward of maxpool3x3s2(relu(bn(x))) given dp = gradient of the pooled output: the pool-input gradient is
    gathered on the fly inside the two BatchNorm backward passes (never written as a tensor).  Returns dx."""
    _chk_bf16(dp, x, dgamma, dbeta)
    N, H, W, Cn = x.shape
    dx = torch.empty_like(x)
    lib = L.lib()
    ws = workspace(lib.kdl_bn_workspace_bytes(N * H * W, Cn), x.device)
    L.check(lib.kdl_bn_bwd_pooled(dp.data_ptr(), idx.data_ptr(), x.data_ptr(), dx.data_ptr(), N, H, W, Cn,
                                  gamma.data_ptr(), beta.data_ptr(), save.data_ptr(), dgamma.data_ptr(),
                                  dbeta.data_ptr(), ws.data_ptr(), ws.numel(), L.stream_ptr()), "kdl_bn_bwd_pooled")
    L.count(2)
    return dx


def maxpool3x3s2_bwd(dy: torch.Tensor, idx: torch.Tensor, in_shape) -> torch.Tensor:
    _chk_bf16(dy)
    N, H, W, Cn = in_shape
    dx = torch.empty(N, H, W, Cn, dtype=torch.bfloat16, device=dy.device)
    L.check(L.lib().kdl_maxpool3x3s2_bwd(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), N, H, W, Cn,
                                         L.stream_ptr()), "kdl_maxpool3x3s2_bwd")
    L.count(1)
    return dx


def avgpool_fwd(x: torch.Tensor) -> torch.Tensor:
    _chk_bf16(x)
    N, H, W, Cn = x.shape
    y = torch.empty(N, Cn, dtype=torch.bfloat16, device=x.device)
    L.check(L.lib().kdl_avgpool_fwd(x.data_ptr(), y.data_ptr(), N, H * W, Cn, L.stream_ptr()),
            "kdl_avgpool_fwd")
    L.count(1)
    return y


def avgpool_bwd(dy: torch.Tensor, in_shape) -> torch.Tensor:
    _chk_bf16(dy)
    N, H, W, Cn = in_shape
    dx = torch.empty(N, H, W, Cn, dtype=torch.bfloat16, device=dy.device)
    L.check(L.lib().kdl_avgpool_bwd(dy.data_ptr(), dx.data_ptr(), N, H * W, Cn, L.stream_ptr()),
            "kdl_avgpool_bwd")
    L.count(1)
    return dx
