"""GBDT boosting-round timing on HIGGS-shaped synthetic data (11M x 28 split over the ranks):
fused histogram+all-reduce kernel vs the same trainer with a separate NCCL all-reduce."""
import json, os, sys, time
sys.path.insert(0, ".")
import torch
import torch.distributed as dist
from kubedl_b200.worker.ddp_trainer import init_distributed
from kubedl_b200.models.gbdt import GBDTTrainer, compute_cuts
from kubedl_b200.ops import lib as L

rank, world, _ = init_distributed("nccl")
dev = torch.device("cuda", torch.cuda.current_device())
TOTAL = int(os.environ.get("ROWS", 11_000_000))
n = TOTAL // world
g = torch.Generator(device=dev).manual_seed(rank)
X = torch.randn(n, 28, device=dev, generator=g)
w = torch.linspace(-1, 1, 28, device=dev)
y = (torch.rand(n, device=dev, generator=g) < torch.sigmoid(X @ w + X[:, 0] * X[:, 1])).float()
cuts = compute_cuts(X)
res = {"world": world, "rows_total": TOTAL, "rows_per_rank": n, "depth": 8}
for comm in (["fused", "nccl"] if world > 1 else ["fused"]):
    tr = GBDTTrainer(X, y, max_depth=8, learning_rate=0.1, cuts=cuts, comm=comm)
    for _ in range(2): tr.grow_tree()
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rounds = 5
    e0.record()
    for _ in range(rounds): tr.grow_tree()
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / rounds], device=dev)
    if world > 1: dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    res[comm] = {"ms_per_round": ms.item(), "rows_per_s": TOTAL / (ms.item() / 1e3), "logloss": float(tr.logloss())}
    del tr
if rank == 0:
    print(json.dumps(res), flush=True)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
