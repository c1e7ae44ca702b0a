"""Stem max-pool kernels in isolation (batch 256, 112x112x64 -> 56x56x64)."""
import sys, json
sys.path.insert(0, ".")
import torch
from kubedl_b200.ops import nn as K
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, iters=8):
    for _ in range(3): fn()
    ts = []
    for _ in range(iters):
        flush.zero_(); torch.cuda._sleep(400000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]
x = torch.randn(256, 112, 112, 64, device=dev).bfloat16()
y, idx = K.maxpool3x3s2_fwd(x)
dy = torch.randn_like(y)
r = {"fwd_ms": timeit(lambda: K.maxpool3x3s2_fwd(x)), "bwd_ms": timeit(lambda: K.maxpool3x3s2_bwd(dy, idx, x.shape))}
r["fwd_bytes_mb"] = (x.numel() * 2 + y.numel() * 3) / 1e6
r["bwd_bytes_mb"] = (x.numel() * 2 + y.numel() * 3) / 1e6
r["fwd_tbs"] = r["fwd_bytes_mb"] / r["fwd_ms"] / 1e3; r["bwd_tbs"] = r["bwd_bytes_mb"] / r["bwd_ms"] / 1e3
print(json.dumps(r))
# BatchNorm-apply + ReLU + max-pool in one pass, as the ResNet stem runs it (finalize kernel + pool kernel)
C = 64
partial = torch.zeros(8, 2, C, device=dev); partial[:, 1] = x.numel() / C / 8      # mean 0, variance 1
gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
save = torch.empty(2, C, device=dev)
fused = lambda: K.bn_relu_maxpool_fwd(x, gamma, beta, None, None, save, (partial, 8))
y2, _ = fused()
ref = torch.relu(x.float())
err = (y2.float() - torch.nn.functional.max_pool2d(ref.permute(0, 3, 1, 2), 3, 2, 1).permute(0, 2, 3, 1)).abs().max().item()
print(json.dumps({"bn_relu_pool_fwd_ms": timeit(fused), "max_abs_err_vs_torch": err}))
