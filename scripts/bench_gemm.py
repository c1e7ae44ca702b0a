"""Per-shape timing of the tcgen05 GEMM vs cuBLAS (torch.matmul) on the ResNet-50 1x1-conv shapes."""
import sys, json, torch
sys.path.insert(0, ".")
from kubedl_b200.ops.gemm import Gemm
dev = torch.device("cuda", 0)
g = Gemm(dev)
B = 256
shapes = [(B*56*56, 64, 64), (B*56*56, 256, 64), (B*56*56, 64, 256), (B*56*56, 128, 256),
          (B*28*28, 512, 128), (B*28*28, 128, 512), (B*14*14, 1024, 256), (B*14*14, 256, 1024),
          (B*7*7, 2048, 512), (B*7*7, 512, 2048), (8192, 8192, 8192)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timeit(fn, iters=10):
    for _ in range(3): fn()
    ts = []
    for _ in range(iters):
        flush.zero_(); torch.cuda._sleep(400000)   # host runs ahead: the bracket times the device only
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts)//2]

rows = []
for (M, N, K) in shapes:
    a = torch.randn(M, K, device=dev).bfloat16(); w = torch.randn(N, K, device=dev).bfloat16()
    dy = torch.randn(M, N, device=dev).bfloat16()
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16); dx = torch.empty(M, K, device=dev, dtype=torch.bfloat16)
    dw = torch.empty(N, K, device=dev, dtype=torch.bfloat16)
    r = {"M": M, "N": N, "K": K}
    r["nt_ours"] = timeit(lambda: g.nt(a, w, out=out)); r["nt_cublas"] = timeit(lambda: torch.matmul(a, w.t(), out=out))
    if M != 8192:
        r["nn_ours"] = timeit(lambda: g.nn(dy, w, out=dx)); r["nn_cublas"] = timeit(lambda: torch.matmul(dy, w, out=dx))
        r["tn_ours"] = timeit(lambda: g.tn(dy, a, out=dw)); r["tn_cublas"] = timeit(lambda: torch.matmul(dy.t(), a, out=dw))
    fl = 2.0 * M * N * K
    by = 2.0 * (M*K + N*K + M*N)
    r["nt_tflops"] = fl / r["nt_ours"] / 1e9; r["nt_gbs"] = by / r["nt_ours"] / 1e6
    r["nt_cublas_tflops"] = fl / r["nt_cublas"] / 1e9
    rows.append(r)
    print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items()}), flush=True)
