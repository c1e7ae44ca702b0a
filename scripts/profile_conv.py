"""One launch of each halo-resident 3x3 convolution kernel on the ResNet-50 layer1 / layer2 shapes
(batch 256) — the program `ncu --set full -k regex:conv3x3_halo` is pointed at."""
import sys
sys.path.insert(0, ".")
import torch
from kubedl_b200.ops import conv
dev = torch.device("cuda", 0)
for (H, C) in [(56, 64), (28, 128)]:
    x = torch.randn(256, H, H, C, device=dev).bfloat16()
    w = (torch.randn(C, 3, 3, C, device=dev) * 0.05).bfloat16()
    for _ in range(2):
        y = conv.fprop(x, w, 1, 1)
        dx = conv.dgrad(y, w, (H, H), 1, 1)
        dw = conv.wgrad(y, x, 3, 3, 1, 1)
    torch.cuda.synchronize()
print("ok")
