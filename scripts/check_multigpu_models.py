"""Multi-GPU check of the GBDT (N3) and DIN (N4/N5) paths under torchrun: fused vs NCCL/oracle."""
import json, os, sys
sys.path.insert(0, ".")
import torch
import torch.distributed as dist
from kubedl_b200.worker.ddp_trainer import init_distributed
from kubedl_b200.models.gbdt import GBDTTrainer, compute_cuts
from kubedl_b200.models.din import DINModel, synthetic_batch

rank, world, _ = init_distributed("nccl")
dev = torch.device("cuda", torch.cuda.current_device())
res = {"world": world}
# ---- GBDT: fused and NCCL variants must grow the same trees; every rank must hold the same tree
n = 300_000
g = torch.Generator(device=dev).manual_seed(rank)
X = torch.randn(n, 28, device=dev, generator=g)
w = torch.linspace(-1, 1, 28, device=dev)
y = (torch.rand(n, device=dev, generator=g) < torch.sigmoid(X @ w + X[:, 0] * X[:, 1])).float()
cuts = compute_cuts(X)
a = GBDTTrainer(X, y, max_depth=6, cuts=cuts, comm="fused")
b = GBDTTrainer(X, y, max_depth=6, cuts=cuts, comm="nccl")
for _ in range(3):
    ta, tb = a.grow_tree(), b.grow_tree()
torch.cuda.synchronize()
same_splits = all(torch.equal(sa[:, :2], sb[:, :2]) for sa, sb in zip(ta.splits, tb.splits))
gathered = [torch.empty_like(ta.leaf_value) for _ in range(world)]
dist.all_gather(gathered, ta.leaf_value)
res["gbdt"] = {"fused_equals_nccl_splits": bool(same_splits), "leaf_max_diff_vs_nccl": float((ta.leaf_value - tb.leaf_value).abs().max()),
               "identical_on_all_ranks": all(torch.equal(gathered[0], t) for t in gathered),
               "logloss_fused": float(a.logloss()), "logloss_nccl": float(b.logloss())}
assert res["gbdt"]["identical_on_all_ranks"] and res["gbdt"]["leaf_max_diff_vs_nccl"] < 1e-3
del a, b
# ---- DIN: sharded table over the ranks vs a single-GPU oracle run of the same global batch
V, L_, B = 80_000, 20, 256
m = DINModel(vocab=V, hist_len=L_, lr=0.1, emb_lr=0.5, seed=1)
full0 = [torch.empty_like(m.shard) for _ in range(world)]
dist.all_gather(full0, m.shard.contiguous())
table0 = m.full_table(full0)
user, cand, hist, label = synthetic_batch(B, L_, V, dev, seed=100 + rank)
loss = m.step(user, cand, hist, label)
torch.cuda.synchronize(); dist.barrier()
full1 = [torch.empty_like(m.shard) for _ in range(world)]
dist.all_gather(full1, m.shard.contiguous())
table1 = m.full_table(full1)
# oracle: gradient of the summed per-rank mean losses w.r.t. the full table
sys.path.insert(0, "tests")
from test_gpu_din import _ref_forward
batches = [None] * world
dist.all_gather_object(batches, (user.cpu(), cand.cpu(), hist.cpu(), label.cpu()))
t = table0.clone().requires_grad_(True)
arena = m.arena
# initial (pre-step) MLP weights are identical on every rank and equal to seed-1 init: rebuild them
ref = DINModel.__new__(DINModel)
from kubedl_b200.models.arena import ParamArena
ar0 = ParamArena()
dims = m.dims
for i, (ki, ko) in enumerate(zip(dims[:-1], dims[1:]), start=1):
    ar0.declare(f"w{i}", (ko, ki), True, "kaiming", ki); ar0.declare(f"b{i}", (ko,), False, "zeros")
ar0.finalize(dev, [n_ for i in (3, 2, 1) for n_ in (f"w{i}", f"b{i}")], heap=None, seed=1)
ar0.w32("w3")[1:].zero_()
mlp = [(ar0.w32(f"w{i}").clone(), ar0.w32(f"b{i}").clone()) for i in (1, 2, 3)]
total = 0
for (u_, c_, h_, l_) in batches:
    lg, _ = _ref_forward(t, mlp, u_.to(dev), c_.to(dev), h_.to(dev))
    total = total + torch.nn.functional.binary_cross_entropy_with_logits(lg, l_.to(dev))
total.backward()
want = -0.5 * t.grad
delta = table1 - table0
touched = want.abs().sum(1) > 0
d, wv = delta[touched].flatten(), want[touched].flatten()
cos = float((d @ wv) / (d.norm() * wv.norm() + 1e-12))
res["din"] = {"loss": float(loss), "rows_touched": int(touched.sum()), "update_cosine_vs_oracle": cos,
              "norm_ratio": float(d.norm() / wv.norm()), "untouched_max_abs": float(delta[~touched].abs().max())}
assert cos > 0.97 and res["din"]["untouched_max_abs"] == 0.0, res["din"]
# dense params stay in lock step across ranks after the fused all-reduce
chk = [torch.empty_like(m.arena.p32) for _ in range(world)]
dist.all_gather(chk, m.arena.p32)
res["din"]["dense_identical_on_all_ranks"] = all(torch.equal(chk[0], c_) for c_ in chk)
assert res["din"]["dense_identical_on_all_ranks"]
# several more steps through the CUDA-graph path: gather -> device barrier -> remote updates -> all-reduce,
# replayed; ranks are deliberately skewed on the host so a missing gather/update order would show
import time
losses = []
for it in range(6):
    if rank == (it % world):
        time.sleep(0.02)
    losses.append(float(m.step(user, cand, hist, label)))
torch.cuda.synchronize(); dist.barrier()
res["din"]["graph_steps_losses"] = [round(x, 5) for x in losses]
res["din"]["status_word"] = int(m.heap.status_word())
assert all(x == x for x in losses) and losses[-1] < losses[0] and res["din"]["status_word"] == 0, res["din"]
chk = [torch.empty_like(m.arena.p32) for _ in range(world)]
dist.all_gather(chk, m.arena.p32)
assert all(torch.equal(chk[0], c_) for c_ in chk)
# ---- generic sharded embedding (dispatch / combine kernels): lookups and updates vs the gathered full table
from kubedl_b200.parallel.embedding import ShardedEmbedding
Ve, De = 40_000, 128
emb = ShardedEmbedding(Ve, De, seed=4)
def full(e):
    parts = [torch.empty_like(e.shard) for _ in range(world)]
    dist.all_gather(parts, e.shard.contiguous())
    return (torch.stack(parts, 1).reshape(-1, De) if e.interleaved else torch.cat(parts, 0))[:Ve]
t0 = full(emb)
ge = torch.Generator(device=dev).manual_seed(77 + rank)
ids_e = (torch.rand(6000, device=dev, generator=ge).pow(3) * Ve).long().clamp_(0, Ve - 1)      # skewed, many duplicates
vec = emb.lookup(ids_e, dtype=torch.float32)
assert torch.equal(vec, t0[ids_e]), "gather mismatch"
emb.barrier()
ge_ = torch.randn(6000, De, device=dev, generator=ge)
emb.apply_grad(ids_e, ge_, lr=0.25)
emb.barrier()
torch.cuda.synchronize(); dist.barrier()
all_ids = [torch.empty_like(ids_e) for _ in range(world)]; dist.all_gather(all_ids, ids_e)
all_g = [torch.empty_like(ge_) for _ in range(world)]; dist.all_gather(all_g, ge_)
want_e = t0.clone()
for i_, g_ in zip(all_ids, all_g):
    want_e.index_add_(0, i_, g_, alpha=-0.25)
err_e = float((full(emb) - want_e).abs().max())
res["sharded_embedding"] = {"rows": Ve, "dim": De, "sharding": emb.sharding, "update_max_err": err_e}
assert err_e < 5e-3, err_e
if rank == 0:
    print(json.dumps(res), flush=True)
dist.barrier(); dist.destroy_process_group()
