"""One launch of each hot kernel on a realistic shape (for `ncu --set full`)."""
import sys
sys.path.insert(0, ".")
import torch
from kubedl_b200.ops import nn as K
from kubedl_b200.ops.gemm import Gemm
dev = torch.device("cuda", 0)
torch.manual_seed(0)
M, C = 256 * 56 * 56, 256
x = torch.randn(M, C, device=dev).bfloat16()
res = torch.randn(M, C, device=dev).bfloat16()
g, b = torch.ones(C, device=dev), torch.zeros(C, device=dev)
rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
for _ in range(2):
    y, save = K.bn_fwd(x, g, b, rm, rv, res=res, relu=True)
    dy = torch.randn_like(y)
    dg, db = torch.empty(C, device=dev, dtype=torch.bfloat16), torch.empty(C, device=dev, dtype=torch.bfloat16)
    dx = K.bn_bwd(dy, x, y, g, save, True, dg, db, dres=dy)
# the variants the model actually uses: statistics from the GEMM epilogue, 1-bit ReLU mask, mask recomputed from x
mask = torch.empty(M * C // 8, dtype=torch.uint8, device=dev)
for _ in range(2):
    y2, _ = K.bn_fwd(x, g, b, None, None, res=res, relu=True, relu_mask=mask)
    dy2 = torch.randn_like(y2)
    dx2 = K.bn_bwd(dy2, x, None, g, save, True, dg, db, dres=dy2, relu_mask=mask)
    dx3 = K.bn_bwd(dy2, x, None, g, save, True, dg, db, beta=b)
gm = Gemm(dev)
a = torch.randn(M, 64, device=dev).bfloat16(); w = torch.randn(256, 64, device=dev).bfloat16()
for _ in range(2):
    o = gm.nt(a, w)            # layer1 conv3 forward: memory-bound
    o2, st = gm.nt_stats(a, w)  # the same with BatchNorm column statistics from the epilogue
    dw = gm.tn(o, a)           # its wgrad (split-K)
A = torch.randn(8192, 8192, device=dev).bfloat16(); B = torch.randn(8192, 8192, device=dev).bfloat16()
for _ in range(2):
    Cc = gm.nt(A, B)           # compute-bound
p32 = torch.randn(25_557_032 // 8 * 8, device=dev); mom = torch.zeros_like(p32)
g16 = torch.randn_like(p32).bfloat16(); p16 = torch.empty_like(g16)
K.sgd_momentum(p32, mom, g16, p16, 0.1, 0.9, 1e-4)
torch.cuda.synchronize()
print("ok")
