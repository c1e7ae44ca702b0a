"""Numerical probe of the halo-resident 3x3 convolution (fprop / dgrad) against conv2d."""
import sys
sys.path.insert(0, ".")
import torch
from kubedl_b200.ops import conv
dev = torch.device("cuda", 0)
for (N, H, C, Co) in [(2, 56, 64, 64), (2, 28, 128, 128), (3, 20, 64, 192), (3, 14, 256, 256), (5, 7, 512, 512), (2, 14, 128, 64)]:
    torch.manual_seed(0)
    x = torch.randn(N, H, H, C, device=dev).bfloat16()
    w = (torch.randn(Co, 3, 3, C, device=dev) * 0.05).bfloat16()
    xr = x.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    wr = w.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    yr = torch.nn.functional.conv2d(xr, wr, stride=1, padding=1)
    y = conv.fprop(x, w, 1, 1)
    torch.cuda.synchronize()
    d = (y.float() - yr.permute(0, 2, 3, 1)).abs()
    print("fprop", N, H, C, Co, "max err", float(d.max()), "ref max", float(yr.abs().max()),
          "bad frac", float((d > 0.05 * yr.abs().max()).float().mean()), flush=True)
    if d.max() > 0.05 * yr.abs().max():
        bad = (d > 0.05 * yr.abs().max()).nonzero()
        print("  first bad idx", bad[:5].tolist(), "rows(h) hist", torch.bincount(bad[:, 1], minlength=H)[:8].tolist(),
              "cols(w) hist", torch.bincount(bad[:, 2], minlength=H)[:8].tolist())
    dy = torch.randn_like(y)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    if conv.halo_supported(1, N, H, H, C, Co):
        dx = conv.dgrad(dy, w, (H, H), 1, 1)
        torch.cuda.synchronize()
        d = (dx.float() - xr.grad.permute(0, 2, 3, 1)).abs()
        print("dgrad", N, H, C, Co, "max err", float(d.max()), "ref max", float(xr.grad.abs().max()), flush=True)
    if True:
        dw = conv.wgrad(dy, x, 3, 3, 1, 1)
        torch.cuda.synchronize()
        ref = wr.grad.permute(0, 2, 3, 1)
        d = (dw.float() - ref).abs()
        print("wgrad", N, H, C, Co, "max err", float(d.max()), "ref max", float(ref.abs().max()),
              "per-tap max err", [round(float(d[:, t // 3, t % 3].max()), 3) for t in range(9)], flush=True)
