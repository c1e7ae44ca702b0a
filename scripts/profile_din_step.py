"""One DIN training step at the BASELINE shape on one GPU, eager (every kernel visible to ncu)."""
import os, sys
sys.path.insert(0, ".")
os.environ["KUBEDL_DIN_GRAPH"] = "0"
import torch
from kubedl_b200.models.din import DINModel, synthetic_batch
dev = torch.device("cuda", 0)
m = DINModel(vocab=10_000_000, hist_len=62, emb_lr=20.0)
b = synthetic_batch(4096, 62, 10_000_000, dev, seed=1)
for _ in range(3):
    m.step(*b)
torch.cuda.synchronize()
torch.cuda.profiler.start()
m.step(*b)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
