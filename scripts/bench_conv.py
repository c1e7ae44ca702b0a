"""3x3 convolution timing: implicit-GEMM tcgen05 kernel vs cuDNN (ResNet-50 shapes, batch 256)."""
import json, sys
sys.path.insert(0, ".")
import torch
from kubedl_b200.ops import conv
dev = torch.device("cuda", 0)
torch.backends.cudnn.benchmark = True
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, iters=8):
    for _ in range(3): fn()
    ts = []
    for _ in range(iters):
        flush.zero_(); torch.cuda._sleep(400000)   # host runs ahead: the bracket times the device only
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]
B = 256
for (H, C, stride) in [(56, 64, 1), (28, 128, 1), (14, 256, 1), (7, 512, 1), (56, 128, 2), (28, 256, 2), (14, 512, 2)]:
    x = torch.randn(B, H, H, C, device=dev).bfloat16(); w = (torch.randn(C, 3, 3, C, device=dev) * 0.05).bfloat16()
    xn, wn = x.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2)
    y = conv.fprop(x, w, stride, 1); dy = torch.randn_like(y); dyn = dy.permute(0, 3, 1, 2)
    r = {"H": H, "C": C, "stride": stride, "gflop": 2 * y.numel() * 9 * C / 1e9}
    r["fprop_ours"] = timeit(lambda: conv.fprop(x, w, stride, 1))
    r["fprop_cudnn"] = timeit(lambda: torch.ops.aten.convolution(xn, wn, None, [stride] * 2, [1, 1], [1, 1], False, [0, 0], 1))
    r["wgrad_ours"] = timeit(lambda: conv.wgrad(dy, x, 3, 3, stride, 1))
    r["wgrad_cudnn"] = timeit(lambda: torch.ops.aten.convolution_backward(dyn, xn, wn, None, [stride] * 2, [1, 1], [1, 1], False, [0, 0], 1, [False, True, False]))
    if stride == 1:
        r["dgrad_ours"] = timeit(lambda: conv.dgrad(dy, w, (H, H), 1, 1))
        r["dgrad_cudnn"] = timeit(lambda: torch.ops.aten.convolution_backward(dyn, xn, wn, None, [1, 1], [1, 1], [1, 1], False, [0, 0], 1, [True, False, False]))
    r["fprop_tflops"] = r["gflop"] / r["fprop_ours"]
    print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items()}), flush=True)
