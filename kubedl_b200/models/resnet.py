"""ResNet-50 (v1.5, torchvision topology) as an explicit forward/backward program over the
kubedl_b200 kernels — no autograd.

Why explicit: every activation is NHWC bf16, every BN(+residual)(+ReLU) is one fused pass,
gradients are written straight into the flat arena that the fused all-reduce works on, the
reducer is told "prefix ready" right after each layer's wgrad, and the whole step (forward,
backward, bucket all-reduces on a side stream, fused SGD) is a static launch sequence that
can be captured once into a CUDA graph.

Convolutions: 1x1 stride-1 convolutions and the FC layer are GEMMs on the [N*H*W, C] view and
go to the tcgen05 GEMM kernel (ops.gemm, BatchNorm statistics from its epilogue); the 7x7 stem runs
straight from the uint8 image (csrc/stem_conv.cu; im2col + GEMM for sizes other than 224x224);
3x3 convolutions go to the halo-resident / TMA-im2col tcgen05 kernels (ops.conv); the 1x1/stride-2
downsample convolutions use the im2col kernel, their dgrad stays a compact GEMM result that the
sibling conv1 dgrad adds in its epilogue; the three stride-2 3x3 dgrads run as four parity
sub-convolutions each on the same im2col kernel — nothing of the step is left on cuDNN.  With gemm=None
everything falls back to cuDNN/cuBLAS (the A/B arm of the tests).
This is config #2 of BASELINE.json ("PyTorchJob ResNet-50 DDP bf16 ... synthetic 224² images").
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from ..ops import conv as conv_ops
from ..ops import nn as K
from .arena import ParamArena

_conv_bwd = torch.ops.aten.convolution_backward


def _nchw(x_nhwc: torch.Tensor) -> torch.Tensor:
    return x_nhwc.permute(0, 3, 1, 2)


def _nhwc(x_nchw: torch.Tensor) -> torch.Tensor:
    y = x_nchw.permute(0, 2, 3, 1)
    return y if y.is_contiguous() else y.contiguous()


class Conv:
    def __init__(self, net: "ResNet50", name: str, cin: int, cout: int, k: int, stride: int, pad: int):
        self.net, self.name = net, name
        self.cin, self.cout, self.k, self.stride, self.pad = cin, cout, k, stride, pad
        net.arena.declare(name, (cout, k, k, cin), decay=True, init="kaiming", fan=cout * k * k)

    def bind(self):
        a = self.net.arena
        self.w = a.w16(self.name)                       # [Cout, kh, kw, Cin] physical
        self.w_nchw = self.w.permute(0, 3, 1, 2)        # channels_last logical view for cuDNN
        self.gw = a.grad(self.name).permute(0, 3, 1, 2)
        self.g_end = a.end_offset(self.name)
        self.is_gemm = (self.k == 1 and self.stride == 1)
        # 3x3 convolutions on the TMA-im2col implicit GEMM (kernel N7) instead of cuDNN
        self.own_conv = (self.k == 3 and self.net.own_conv and self.net.gemm is not None
                         and self.cin % 64 == 0)
        # 1x1 / stride-2 downsample convolution: fprop and wgrad on the TMA-im2col kernel, dgrad as a
        # plain GEMM whose COMPACT result the sibling conv1 dgrad adds at even pixels in its epilogue
        self.own_strided = (self.k == 1 and self.stride == 2 and self.net.own_conv and self.net.gemm is not None
                            and self.cin % 64 == 0)

    def fwd(self, x: torch.Tensor, want_stats: bool = False):
        """want_stats: also return the BN column statistics produced by the GEMM epilogue (None on
        the cuDNN path)."""
        net = self.net
        if self.is_gemm and net.gemm is not None:
            N, H, W, Cin = x.shape
            if want_stats and net.fuse_bn_stats:
                y, stats = net.gemm.nt_stats(x.view(-1, Cin), self.w.view(self.cout, Cin))
                return y.view(N, H, W, self.cout), stats
            y = net.gemm.nt(x.view(-1, Cin), self.w.view(self.cout, Cin)).view(N, H, W, self.cout)
            return (y, None) if want_stats else y
        if self.own_conv or self.own_strided:
            if want_stats and net.fuse_bn_stats and net.fuse_conv_stats:   # im2col kernel: statistics from its epilogue
                return conv_ops.fprop(x, self.w, self.stride, self.pad, want_stats=True)
            y = conv_ops.fprop(x, self.w, self.stride, self.pad)
            return (y, None) if want_stats else y
        y = torch.ops.aten.convolution(_nchw(x), self.w_nchw, None, [self.stride] * 2,
                                       [self.pad] * 2, [1, 1], False, [0, 0], 1)
        return (_nhwc(y), None) if want_stats else _nhwc(y)

    def bwd(self, dy: torch.Tensor, x: torch.Tensor, need_dx: bool = True,
            addend: Optional[torch.Tensor] = None, addend_compact: bool = False,
            addend_mask: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
        """Returns dx (+ addend, the gradient arriving over the residual branch — fused into the
        dgrad GEMM epilogue on the tcgen05 path; addend_compact: it is the [N, H/2, W/2, Cin] result
        of a stride-2 sibling's bwd_compact; addend_mask: the addend still has to pass the block's output
        ReLU, whose 1-bit mask the epilogue applies — GEMM path only)."""
        net = self.net
        if self.is_gemm and net.gemm is not None:
            N, H, W, Cin = x.shape
            dy2, x2 = dy.view(-1, self.cout), x.view(-1, Cin)
            # dW[Cout, Cin] = dy^T x ; dx[M, Cin] = dy W (+ addend)
            net.gemm.tn(dy2, x2, out=net.arena.grad(self.name).view(self.cout, Cin))
            dx = None
            if need_dx:
                dx = net.gemm.nn(dy2, self.w.view(self.cout, Cin),
                                 addend=None if addend is None else addend.view(-1, Cin),
                                 addend_hw=(H, W) if addend_compact else None,
                                 addend_mask=addend_mask).view(N, H, W, Cin)
        elif self.own_conv:
            conv_ops.wgrad(dy, x, self.k, self.k, self.stride, self.pad, out=net.arena.grad(self.name))
            dx = None
            if need_dx and self.stride == 1:
                dx = conv_ops.dgrad(dy, self.w, x.shape[1:3], 1, self.pad, addend=addend)
            elif need_dx and net.own_strided_dgrad and conv_ops.dgrad_s2_supported(
                    x.shape[1:3], self.cin, self.cout, self.k, self.k, self.stride, self.pad):
                dx = conv_ops.dgrad_s2(dy, self.w, x.shape[1:3])       # four parity sub-convolutions
                if addend is not None:
                    K.add_(dx, addend)
            elif need_dx:     # other strided shapes: cuDNN
                dxn, _, _ = _conv_bwd(_nchw(dy), _nchw(x), self.w_nchw, None, [self.stride] * 2,
                                      [self.pad] * 2, [1, 1], False, [0, 0], 1, [True, False, False])
                dx = _nhwc(dxn)
                if addend is not None:
                    K.add_(dx, addend)
        else:
            dxn, dw, _ = _conv_bwd(_nchw(dy), _nchw(x), self.w_nchw, None, [self.stride] * 2,
                                   [self.pad] * 2, [1, 1], False, [0, 0], 1, [need_dx, True, False])
            self.gw.copy_(dw)
            dx = _nhwc(dxn) if need_dx else None
            if dx is not None and addend is not None:
                K.add_(dx, addend)
        net.reducer_ready(self.g_end)
        return dx

    def bwd_compact(self, dy: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
        """1x1 / stride-2: weight gradient into the arena, and the input gradient at the sampled (even)
        pixels only, [N, H/2, W/2, Cin]."""
        net = self.net
        conv_ops.wgrad(dy, x, 1, 1, self.stride, 0, out=net.arena.grad(self.name))
        dxc = net.gemm.nn(dy.view(-1, self.cout), self.w.view(self.cout, self.cin))
        net.reducer_ready(self.g_end)
        return dxc.view(dy.shape[0], dy.shape[1], dy.shape[2], self.cin)


class BN:
    def __init__(self, net: "ResNet50", name: str, c: int, relu: bool, zero_init: bool = False):
        self.net, self.name, self.c, self.relu = net, name, c, relu
        net.arena.declare(name + ".weight", (c,), decay=False, init="zeros" if zero_init else "ones")
        net.arena.declare(name + ".bias", (c,), decay=False, init="zeros")

    def bind(self):
        a, dev = self.net.arena, self.net.device
        self.gamma, self.beta = a.w32(self.name + ".weight"), a.w32(self.name + ".bias")
        self.dgamma, self.dbeta = a.grad(self.name + ".weight"), a.grad(self.name + ".bias")
        self.running_mean = torch.zeros(self.c, dtype=torch.float32, device=dev)
        self.running_var = torch.ones(self.c, dtype=torch.float32, device=dev)
        self.save = torch.empty(2, self.c, dtype=torch.float32, device=dev)

    def fwd(self, x, res=None, stats=None, train=True):
        # ReLU after a residual add: the mask cannot be recomputed from x alone, so the forward pass
        # leaves one bit per element for the backward pass (instead of the backward reading y twice)
        self.mask = None
        if res is not None and self.relu and train and self.net.bn_relu_bitmask:
            self.mask = torch.empty(x.numel() // 8, dtype=torch.uint8, device=x.device)
        y, _ = K.bn_fwd(x, self.gamma, self.beta, self.running_mean, self.running_var, res=res,
                        relu=self.relu, save=self.save, stats=stats, relu_mask=self.mask)
        return y

    def bwd(self, dy, x, y, dres=None, had_residual=False):
        """y is read for the ReLU mask only when the forward pass added a residual; otherwise the mask
        is recomputed from x (one tensor less to stream in both backward passes)."""
        mask = getattr(self, "mask", None) if had_residual else None
        self.mask = None
        return K.bn_bwd(dy, x, y if (self.relu and had_residual and mask is None) else None, self.gamma, self.save,
                        self.relu, self.dgamma, self.dbeta, dres=dres, beta=self.beta, relu_mask=mask)


class Bottleneck:
    def __init__(self, net, name, cin, width, stride, downsample):
        cout = width * 4
        self.c1 = Conv(net, f"{name}.conv1", cin, width, 1, 1, 0)
        self.b1 = BN(net, f"{name}.bn1", width, relu=True)
        self.c2 = Conv(net, f"{name}.conv2", width, width, 3, stride, 1)
        self.b2 = BN(net, f"{name}.bn2", width, relu=True)
        self.c3 = Conv(net, f"{name}.conv3", width, cout, 1, 1, 0)
        self.b3 = BN(net, f"{name}.bn3", cout, relu=True, zero_init=net.zero_init_residual)
        self.cd = self.bd = None
        if downsample:
            self.cd = Conv(net, f"{name}.downsample.0", cin, cout, 1, stride, 0)
            self.bd = BN(net, f"{name}.downsample.1", cout, relu=False)
        self.saved = None

    def modules(self):
        ms = [self.c1, self.b1, self.c2, self.b2, self.c3, self.b3]
        if self.cd is not None:
            ms += [self.cd, self.bd]
        return ms

    def backward_param_order(self) -> List[str]:
        # must match the order in which bwd() produces the weight gradients: conv3, conv2,
        # downsample (if any), conv1 — the reducer treats the arena as a growing ready prefix
        names = [self.c3.name, self.b3.name + ".weight", self.b3.name + ".bias",
                 self.c2.name, self.b2.name + ".weight", self.b2.name + ".bias"]
        if self.cd is not None:
            names += [self.cd.name, self.bd.name + ".weight", self.bd.name + ".bias"]
        names += [self.c1.name, self.b1.name + ".weight", self.b1.name + ".bias"]
        return names

    def fwd(self, x, train=True):
        c1, s1 = self.c1.fwd(x, want_stats=True)
        a1 = self.b1.fwd(c1, stats=s1)
        c2, s2 = self.c2.fwd(a1, want_stats=True)
        a2 = self.b2.fwd(c2, stats=s2)
        c3, s3 = self.c3.fwd(a2, want_stats=True)
        if self.cd is not None:
            cdo, sd = self.cd.fwd(x, want_stats=True)
            idn = self.bd.fwd(cdo, stats=sd)
        else:
            cdo, idn = None, x
        out = self.b3.fwd(c3, res=idn, stats=s3, train=train)
        if train:
            self.saved = (x, c1, a1, c2, a2, c3, cdo, out)
        return out

    def bwd(self, dout, need_dx=True):
        x, c1, a1, c2, a2, c3, cdo, out = self.saved
        self.saved = None
        # Identity blocks: dz = dout * relu'(out) is needed once more, as the residual term of this block's
        # input gradient.  Instead of having the BatchNorm backward write it (one more activation-sized
        # store), conv1's dgrad epilogue reads dout and the 1-bit mask itself.
        net = self.c1.net
        mask_addend = (self.cd is None and net.mask_addend and self.b3.mask is not None
                       and self.c1.is_gemm and net.gemm is not None)
        amask = self.b3.mask if mask_addend else None
        dc3 = self.b3.bwd(dout, c3, out, dres=None if mask_addend else dout, had_residual=True)
        dres = dout       # dz unless mask_addend (then still the unmasked dout, masked by the consumer)
        da2 = self.c3.bwd(dc3, a2)
        dc2 = self.b2.bwd(da2, c2, a2)
        da1 = self.c2.bwd(dc2, a1)
        dc1 = self.b1.bwd(da1, c1, a1)
        if self.cd is not None:
            dcd = self.bd.bwd(dres, cdo, None)
            if self.cd.own_strided and self.c1.is_gemm:
                dx = self.c1.bwd(dc1, x, addend=self.cd.bwd_compact(dcd, x), addend_compact=True)
            else:
                dx = self.c1.bwd(dc1, x, addend=self.cd.bwd(dcd, x))
        else:
            dx = self.c1.bwd(dc1, x, addend=dres, addend_mask=amask)
        return dx


class ResNet50:
    LAYERS = [(64, 3, 1), (128, 4, 2), (256, 6, 2), (512, 3, 2)]

    def __init__(self, device="cuda", num_classes: int = 1000, heap=None, gemm=None, seed: int = 0,
                 zero_init_residual: bool = False):
        self.device = torch.device(device)
        self.zero_init_residual = zero_init_residual
        import os as _os
        # BN column statistics from the GEMM epilogue (1x1 convolutions, stem): every epilogue warp keeps
        # the sums of its columns in registers across the persistent loop, a small kernel reduces the
        # per-warp rows, and the statistics pass over the activation disappears: 22.0 vs 22.5 ms/step
        # (profiles/bench_1gpu_bnstats_ab.log).  The first version used shared-memory float atomics
        # (CAS loops) and a serial finalizer and was slower than the separate pass.
        self.fuse_bn_stats = _os.environ.get("KUBEDL_FUSE_BN_STATS", "1") == "1"
        self.fuse_conv_stats = _os.environ.get("KUBEDL_FUSE_CONV_STATS", "1") == "1"
        # ReLU masks of the residual BatchNorms as 1 bit per element (0: the backward pass reads y)
        self.bn_relu_bitmask = _os.environ.get("KUBEDL_BN_BITMASK", "1") == "1"
        # identity blocks: the residual gradient is masked inside conv1's dgrad epilogue (no dz tensor)
        self.mask_addend = _os.environ.get("KUBEDL_MASK_ADDEND", "1") == "1"
        # stem convolution straight from the uint8 image (224x224 inputs); 0: im2col patch matrix + GEMM
        self.stem_direct = _os.environ.get("KUBEDL_STEM_DIRECT", "1") == "1"
        # the stem as a 4x4 convolution on a space-to-depth copy of the image (csrc/stem_s2d.cu): operand tiles come
        # from TMA instead of being built by threads.  Pack 0.03 + forward 0.19 + wgrad 0.17 ms against 0.33 + 0.30 for
        # the direct kernels; 18.85 -> 18.69 ms/step on the same box.  KUBEDL_STEM_S2D=0 selects the direct kernels.
        self.stem_s2d = _os.environ.get("KUBEDL_STEM_S2D", "1") == "1"
        self._s2d_buf = None
        self.fuse_stem_pool = _os.environ.get("KUBEDL_FUSE_STEM_POOL", "1") == "1"
        self.fuse_stem_pool_bwd = _os.environ.get("KUBEDL_FUSE_STEM_POOL_BWD", "0") == "1"
        # 3x3 convolutions on our own tcgen05 kernels (N7: halo-resident kernels for stride 1, TMA-im2col
        # implicit GEMM for stride 2) instead of cuDNN.  End to end the two are within 1 % of each other
        # (23.1 vs 22.9 ms/step, profiles/bench_1gpu_ownconv_halo.log); KUBEDL_OWN_CONV=0 restores cuDNN.
        self.own_conv = _os.environ.get("KUBEDL_OWN_CONV", "1") == "1"
        # the three 3x3 / stride-2 input gradients on our im2col kernel too (0: cuDNN)
        self.own_strided_dgrad = _os.environ.get("KUBEDL_OWN_STRIDED_DGRAD", "1") == "1"
        self.num_classes = num_classes
        self.arena = ParamArena()
        self.gemm = gemm
        self.reducer = None
        self.stem = Conv(self, "conv1", 3, 64, 7, 2, 3)
        self.stem_bn = BN(self, "bn1", 64, relu=True)
        self.blocks: List[Bottleneck] = []
        cin = 64
        for li, (width, n, stride) in enumerate(self.LAYERS, start=1):
            for bi in range(n):
                s = stride if bi == 0 else 1
                self.blocks.append(Bottleneck(self, f"layer{li}.{bi}", cin, width, s, bi == 0))
                cin = width * 4
        self.fc_in = cin
        self.arena.declare("fc.weight", (num_classes, cin), decay=True, init="fc", fan=cin)
        self.arena.declare("fc.bias", (num_classes,), decay=False, init="fc", fan=cin)
        order = ["fc.weight", "fc.bias"]
        for b in reversed(self.blocks):
            order += b.backward_param_order()
        order += ["conv1", "bn1.weight", "bn1.bias"]
        self.arena.finalize(self.device, order, heap=heap, seed=seed)
        for m in [self.stem, self.stem_bn] + [m for b in self.blocks for m in b.modules()]:
            m.bind()
        self.fc_w = self.arena.w16("fc.weight")
        self.fc_b = self.arena.w16("fc.bias")
        self.saved = None

    @property
    def num_params(self) -> int:
        return sum(s.numel for s in self.arena.layout)

    def reducer_ready(self, end_elem: int) -> None:
        if self.reducer is not None:
            self.reducer.ready_upto(end_elem)

    # ------------------------------------------------------------------ forward
    def _stem_w_padded(self) -> torch.Tensor:
        return K.stem_weight_padded(self.stem.w)

    def forward(self, x, train: bool = True) -> torch.Tensor:
        """x: normalised bf16 NHWC image, or (tcgen05 path) the raw uint8 NHWC image — then the stem
        is im2col (fused with the normalisation) + GEMM instead of cuDNN's 3-channel convolution."""
        stem_gemm = x.dtype == torch.uint8
        if stem_gemm:
            assert self.gemm is not None
            N, H, W, _ = x.shape
            if self.stem_s2d and K.stem_conv_supported(H, W):
                if self._s2d_buf is None or self._s2d_buf.shape[0] != N:
                    self._s2d_buf = torch.empty(N, 115, 115, 16, dtype=torch.bfloat16, device=x.device)
                s2d = K.stem_s2d_pack(x, self._s2d_buf)
                c, stem_stats = K.stem_s2d_fwd(s2d, self._stem_w_padded(), want_stats=self.fuse_bn_stats)
                x_bf16_nhwc = s2d
            elif self.stem_direct and K.stem_conv_supported(H, W):
                # straight from the uint8 image: patch tiles are built in shared memory (csrc/stem_conv.cu)
                c, stem_stats = K.stem_conv_fwd(x, self._stem_w_padded(), want_stats=self.fuse_bn_stats)
                x_bf16_nhwc = x
            else:
                patches = K.stem_im2col(x)
                if self.fuse_bn_stats:
                    c, stem_stats = self.gemm.nt_stats(patches, self._stem_w_padded())
                else:
                    c, stem_stats = self.gemm.nt(patches, self._stem_w_padded()), None
                c = c.view(N, (H - 1) // 2 + 1, (W - 1) // 2 + 1, 64)
                x_bf16_nhwc = patches
        else:
            x_bf16_nhwc = x
            c, stem_stats = self.stem.fwd(x_bf16_nhwc), None
        if stem_stats is not None and self.fuse_stem_pool:
            # BN apply + ReLU + max-pool in one pass over the convolution output; the normalised
            # activation is not needed by the backward pass either (its ReLU mask is recomputed from c)
            sb = self.stem_bn
            p, idx = K.bn_relu_maxpool_fwd(c, sb.gamma, sb.beta, sb.running_mean, sb.running_var, sb.save, stem_stats)
            a = None
        else:
            a = self.stem_bn.fwd(c, stats=stem_stats)
            p, idx = K.maxpool3x3s2_fwd(a)
        h = p
        for b in self.blocks:
            h = b.fwd(h, train)
        pooled = K.avgpool_fwd(h)
        if self.gemm is not None:
            logits = self.gemm.nt(pooled, self.fc_w, bias=self.fc_b)
        else:
            logits = torch.addmm(self.fc_b, pooled, self.fc_w.t())
        if train:
            self.saved = (x_bf16_nhwc, c, a, idx, h.shape, pooled)
        return logits

    # ------------------------------------------------------------------ backward
    def backward(self, dlogits: torch.Tensor) -> None:
        x, c, a, idx, hshape, pooled = self.saved
        self.saved = None
        if self.reducer is not None:
            self.reducer.start()
        a_ = self.arena
        gfc = a_.grad("fc.weight")
        if self.gemm is not None:
            self.gemm.tn(dlogits, pooled, out=gfc)
            dpooled = self.gemm.nn(dlogits, self.fc_w)
        else:
            torch.mm(dlogits.t(), pooled, out=gfc)
            dpooled = torch.mm(dlogits, self.fc_w)
        torch.sum(dlogits, dim=0, out=a_.grad("fc.bias"))
        self.reducer_ready(a_.end_offset("fc.weight"))
        dh = K.avgpool_bwd(dpooled, hshape)
        for b in reversed(self.blocks):
            dh = b.bwd(dh)
        if self.fuse_stem_pool_bwd:
            # max-pool backward gathered inside the two BatchNorm backward passes: the 411 MB pool-input
            # gradient is never written or re-read.  Bit-identical, but measured SLOWER (+0.8 ms/step: the
            # gather costs more than the traffic it saves, profiles/stem_pool_ab.log), so off by default.
            sb = self.stem_bn
            dc = K.bn_bwd_pooled(dh, idx, c, sb.gamma, sb.beta, sb.save, sb.dgamma, sb.dbeta)
        else:
            da = K.maxpool3x3s2_bwd(dh, idx, c.shape)
            dc = self.stem_bn.bwd(da, c, a)
        if x.dim() == 4 and x.shape[1:] == (115, 115, 16):     # space-to-depth stem: the same TMA tiles as the forward
            gw = K.stem_s2d_wgrad(x, dc)
            self.arena.grad("conv1").view(64, 147).copy_(K.stem_weight_unpad(gw))
            self.reducer_ready(self.stem.g_end)
        elif x.dtype == torch.uint8:    # direct stem kernels: patch tiles rebuilt in shared memory
            gw = K.stem_conv_wgrad(x, dc)
            self.arena.grad("conv1").view(64, 147).copy_(K.stem_weight_unpad(gw))
            self.reducer_ready(self.stem.g_end)
        elif x.dim() == 2:   # im2col patches: stem wgrad is one tn GEMM (split over the 3.2 M rows)
            gw = self.gemm.tn(dc.view(-1, 64), x)
            self.arena.grad("conv1").view(64, 147).copy_(K.stem_weight_unpad(gw))
            self.reducer_ready(self.stem.g_end)
        else:
            self.stem.bwd(dc, x, need_dx=False)
        if self.reducer is not None:
            self.reducer.finish()

    def loss_and_grad(self, images_u8: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        x = images_u8 if self.gemm is not None else K.u8_normalize(images_u8, cpad=3)
        logits = self.forward(x, train=True)
        loss, dlogits = K.softmax_xent(logits, labels, V=self.num_classes)
        self.backward(dlogits)
        return loss

    def sgd_step(self, lr: float, momentum: float = 0.9, weight_decay: float = 1e-4) -> None:
        a = self.arena
        K.sgd_momentum(a.p32, a.mom, a.g16, a.p16, lr, momentum, weight_decay, wd_count=a.wd_count)
