"""Two boosting rounds on 11M x 28 (1 GPU) — the program the ncu launch list is taken from."""
import sys
sys.path.insert(0, ".")
import torch
from kubedl_b200.models.gbdt import GBDTTrainer, compute_cuts
dev = torch.device("cuda", 0)
n = 11_000_000
g = torch.Generator(device=dev).manual_seed(0)
X = torch.randn(n, 28, device=dev, generator=g)
w = torch.linspace(-1, 1, 28, device=dev)
y = (torch.rand(n, device=dev, generator=g) < torch.sigmoid(X @ w + X[:, 0] * X[:, 1])).float()
tr = GBDTTrainer(X, y, max_depth=8, learning_rate=0.1, cuts=compute_cuts(X))
tr.grow_tree()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
tr.grow_tree()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("ok", float(tr.logloss()))
