"""One launch of each kernel changed in round 2, on the ResNet-50 layer-1 shapes (for `ncu --set full`)."""
import sys
sys.path.insert(0, ".")
import torch
from kubedl_b200.ops import nn as K
from kubedl_b200.ops.gemm import Gemm

dev = torch.device("cuda", 0)
torch.manual_seed(0)
M, C = 256 * 56 * 56, 256
gm = Gemm(dev)
x = torch.randn(M, C, device=dev).bfloat16()
res = torch.randn(M, C, device=dev).bfloat16()
g, b = torch.ones(C, device=dev), torch.zeros(C, device=dev)
mask = torch.empty(M * C // 8, dtype=torch.uint8, device=dev)
dg = torch.empty(C, device=dev, dtype=torch.bfloat16)
db = torch.empty(C, device=dev, dtype=torch.bfloat16)
dc1 = torch.randn(M, 64, device=dev).bfloat16()
w1 = (torch.randn(64, C, device=dev) * 0.05).bfloat16()
a2 = torch.randn(M, 64, device=dev).bfloat16()
w3 = (torch.randn(C, 64, device=dev) * 0.05).bfloat16()
for _ in range(2):
    y, save = K.bn_fwd(x, g, b, None, None, res=res, relu=True, relu_mask=mask)
    dout = torch.randn_like(y)
    dc3 = K.bn_bwd(dout, x, None, g, save, True, dg, db, relu_mask=mask)      # bn_reduce<1,3> + bn_dx<.,3> (no dz write)
    dx = gm.nn(dc1, w1, addend=dout, addend_mask=mask)                        # conv1 dgrad + masked residual addend (8 epilogue warps)
    c3, st = gm.nt_stats(a2, w3)                                              # conv3 forward + BN statistics (8 epilogue warps)
torch.cuda.synchronize()
print("ok")
