"""Implicit-GEMM convolution (TMA im2col + tcgen05, kernel N7) vs torch conv2d in fp32."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cos(a, b):
    a, b = a.flatten().float(), b.flatten().float()
    return float((a @ b) / (a.norm() * b.norm() + 1e-12))


@pytest.mark.parametrize("N,H,Cin,Cout,R,stride,pad", [
    (8, 56, 64, 64, 3, 1, 1), (4, 28, 128, 128, 3, 1, 1), (3, 14, 256, 256, 3, 1, 1),
    (4, 56, 128, 128, 3, 2, 1), (2, 14, 512, 512, 3, 1, 1), (5, 20, 64, 192, 3, 1, 1),
    (4, 56, 256, 512, 1, 2, 0), (3, 14, 1024, 2048, 1, 2, 0),       # the 1x1 / stride-2 downsample convolutions
])
@pytest.mark.parametrize("halo", [True, False])
def test_conv_fprop_dgrad_wgrad(cuda, N, H, Cin, Cout, R, stride, pad, halo):
    """halo=True lets the 3x3/s1 shapes that fit take the halo-resident kernel (conv3x3_halo.cu);
    halo=False pins everything on the TMA-im2col kernel."""
    from kubedl_b200.ops import conv
    torch.manual_seed(0)
    x = torch.randn(N, H, H, Cin, device=cuda).bfloat16()
    w = (torch.randn(Cout, R, R, Cin, device=cuda) * 0.05).bfloat16()
    xr = x.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    wr = w.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    yr = torch.nn.functional.conv2d(xr, wr, stride=stride, padding=pad)
    y = conv.fprop(x, w, stride, pad, halo=halo)
    assert y.shape == yr.permute(0, 2, 3, 1).shape
    err = float((y.float() - yr.permute(0, 2, 3, 1)).abs().max() / yr.abs().max())
    assert err < 1.5e-2, err
    dy = torch.randn_like(y)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    dw = conv.wgrad(dy, x, R, R, stride, pad, halo=halo)
    assert _cos(dw, wr.grad.permute(0, 2, 3, 1)) > 0.999
    err = float((dw.float() - wr.grad.permute(0, 2, 3, 1)).abs().max() / wr.grad.abs().max())
    assert err < 2e-2, err
    if stride == 1:
        dx = conv.dgrad(dy, w, (H, H), stride, pad, halo=halo)
        err = float((dx.float() - xr.grad.permute(0, 2, 3, 1)).abs().max() / xr.grad.abs().max())
        assert err < 1.5e-2, err
        add = torch.randn_like(dx)
        dx2 = conv.dgrad(dy, w, (H, H), stride, pad, addend=add, halo=halo)
        err = float((dx2.float() - (xr.grad.permute(0, 2, 3, 1) + add.float())).abs().max() / xr.grad.abs().max())
        assert err < 2e-2, err


def test_halo_dispatch(cuda):
    from kubedl_b200.ops import conv
    assert conv.halo_supported(0, 256, 56, 56, 64, 64) and conv.halo_supported(1, 256, 28, 28, 128, 128)
    assert not conv.halo_supported(0, 256, 14, 14, 256, 256)     # weights do not fit: im2col kernel


@pytest.mark.parametrize("N,H,W,Cin,Cout", [
    (2, 8, 8, 64, 64),          # 1-CTA kernel, BN = 64
    (3, 14, 14, 128, 128),      # CTA pair, 128-wide tiles, odd H/2
    (4, 28, 28, 256, 256),      # CTA pair, 256-wide tiles (ResNet layer3.0)
    (16, 56, 56, 128, 128),     # ResNet layer2.0 at a smaller batch
    (5, 14, 14, 512, 512),      # ResNet layer4.0
])
def test_conv3x3_stride2_dgrad_matches_reference(cuda, N, H, W, Cin, Cout):
    """3x3 / stride-2 / pad-1 input gradient as four parity sub-convolutions == conv_transpose in fp32."""
    from kubedl_b200.ops import conv as conv_ops
    torch.manual_seed(31)
    dy = torch.randn(N, H // 2, W // 2, Cout, device=cuda).bfloat16()
    w = (torch.randn(Cout, 3, 3, Cin, device=cuda) * 0.05).bfloat16()
    dx = conv_ops.dgrad_s2(dy, w, (H, W))
    ref = torch.nn.functional.conv_transpose2d(dy.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2),
                                               stride=2, padding=1, output_padding=1).permute(0, 2, 3, 1)
    assert ref.shape == dx.shape
    err = float((dx.float() - ref).abs().max() / ref.abs().max())
    assert err < 1.5e-2, err
    # every parity class was written (no pixel left at its uninitialised value)
    for ph in (0, 1):
        for pw in (0, 1):
            sub, rsub = dx[:, ph::2, pw::2].float(), ref[:, ph::2, pw::2]
            assert float((sub - rsub).abs().max() / ref.abs().max()) < 1.5e-2, (ph, pw)


@pytest.mark.parametrize("N,H,W,Cin,Cout,R,stride,pad", [
    (8, 14, 14, 256, 256, 3, 1, 1),      # layer3 3x3 (CTA pair, 256-wide tiles)
    (6, 28, 28, 256, 256, 3, 2, 1),      # layer3.0 3x3 / stride 2
    (4, 56, 56, 256, 512, 1, 2, 0),      # layer2.0 downsample 1x1 / stride 2
    (3, 8, 8, 64, 64, 3, 2, 1),          # 1-CTA kernel, BN = 64
])
def test_conv_fprop_fused_bn_statistics(cuda, N, H, W, Cin, Cout, R, stride, pad):
    """Column statistics from the im2col convolution's epilogue == statistics of the stored output, and the
    BatchNorm fed by them == the BatchNorm that recomputes them."""
    from kubedl_b200.ops import conv as conv_ops
    from kubedl_b200.ops import nn as Kn
    torch.manual_seed(41)
    x = torch.randn(N, H, W, Cin, device=cuda).bfloat16()
    w = (torch.randn(Cout, R, R, Cin, device=cuda) * 0.05).bfloat16()
    y, stats = conv_ops.fprop(x, w, stride, pad, halo=False, want_stats=True)
    assert stats is not None
    partial, nparts = stats
    y0 = conv_ops.fprop(x, w, stride, pad, halo=False)
    assert torch.equal(y, y0)
    tot = partial[:nparts].sum(0)
    yf = y.float().view(-1, Cout)
    torch.testing.assert_close(tot[0], yf.sum(0), atol=2e-1, rtol=2e-3)
    torch.testing.assert_close(tot[1], (yf * yf).sum(0), atol=2e-1, rtol=2e-3)
    gamma, beta = torch.rand(Cout, device=cuda) + 0.5, torch.randn(Cout, device=cuda)
    o1, sv1 = Kn.bn_fwd(y, gamma, beta, None, None, relu=True)
    o2, sv2 = Kn.bn_fwd(y, gamma, beta, None, None, relu=True, stats=stats)
    torch.testing.assert_close(sv2, sv1, atol=1e-4, rtol=1e-4)
    torch.testing.assert_close(o2.float(), o1.float(), atol=2e-2, rtol=2e-2)
