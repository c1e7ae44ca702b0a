"""One launch of the direct stem kernels + the stem BN / pool kernels at batch 256 (ncu target)."""
import sys
sys.path.insert(0, ".")
import torch
from kubedl_b200.ops import nn as K
dev = torch.device("cuda", 0)
img = torch.randint(0, 256, (256, 224, 224, 3), dtype=torch.uint8, device=dev)
w = (torch.randn(64, 7, 7, 3, device=dev) * 0.05).bfloat16()
wp = K.stem_weight_padded(w)
for _ in range(2):
    y, stats = K.stem_conv_fwd(img, wp, want_stats=True)
    gw = K.stem_conv_wgrad(img, y)
torch.cuda.synchronize()
print("ok")
# device time per launch (CUDA events, L2 flushed between iterations)
import json
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, iters=8):
    ts = []
    for _ in range(iters):
        flush.zero_(); torch.cuda._sleep(400000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]
print(json.dumps({"stem_fprop_stats_ms": timeit(lambda: K.stem_conv_fwd(img, wp, want_stats=True)),
                  "stem_wgrad_ms": timeit(lambda: K.stem_conv_wgrad(img, y))}))
# the same layer on the space-to-depth image (csrc/stem_s2d.cu)
s2d = K.stem_s2d_pack(img)
print(json.dumps({"s2d_pack_ms": timeit(lambda: K.stem_s2d_pack(img, s2d)),
                  "s2d_fprop_stats_ms": timeit(lambda: K.stem_s2d_fwd(s2d, wp, want_stats=True)),
                  "s2d_wgrad_ms": timeit(lambda: K.stem_s2d_wgrad(s2d, y))}))
