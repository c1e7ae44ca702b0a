"""BatchNorm backward on the shapes that dominate the ResNet-50 step: event-timed (L2 flushed between
iterations) and, under `ncu --set full -k regex:bn_`, one profiled launch of each variant."""
import sys
sys.path.insert(0, ".")
import torch
from kubedl_b200.ops import nn as K

dev = torch.device("cuda", 0)
torch.manual_seed(0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


for (M, C) in [(256 * 56 * 56, 256), (256 * 56 * 56, 64), (256 * 28 * 28, 512), (256 * 14 * 14, 1024)]:
    x = torch.randn(M, C, device=dev).bfloat16()
    res = torch.randn(M, C, device=dev).bfloat16()
    g, b = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    mask = torch.empty(M * C // 8, dtype=torch.uint8, device=dev)
    y, save = K.bn_fwd(x, g, b, None, None, res=res, relu=True, relu_mask=mask)
    dy = torch.randn_like(y)
    dg = torch.empty(C, device=dev, dtype=torch.bfloat16)
    db = torch.empty(C, device=dev, dtype=torch.bfloat16)
    dx = torch.empty_like(x)
    mb = M * C * 2 / 1e6
    t_mask_dres = timed(lambda: K.bn_bwd(dy, x, None, g, save, True, dg, db, dx=dx, dres=dy, relu_mask=mask))
    t_mask = timed(lambda: K.bn_bwd(dy, x, None, g, save, True, dg, db, dx=dx, relu_mask=mask))
    t_recomp = timed(lambda: K.bn_bwd(dy, x, None, g, save, True, dg, db, dx=dx, beta=b))
    print(f"M={M} C={C} tensor={mb:.0f}MB  bwd(mask,+dz) {t_mask_dres:.1f}us  bwd(mask) {t_mask:.1f}us "
          f"[{(5 * mb + mb / 8) / t_mask:.0f} GB/s... bytes 5.125T]  bwd(recompute) {t_recomp:.1f}us "
          f"[{5 * mb / t_recomp / 1e3:.2f} TB/s]", flush=True)
    del x, res, y, dy, dx, mask
torch.cuda.synchronize()
print("ok")
