"""DIN step timing at the BASELINE config (10M-row table sharded over the ranks, B=4096/GPU,
62-item histories): fused NVLink gather / scatter-update kernels vs a stock implementation
(all_to_all_single id+vector exchange, index_select, autograd attention/MLP, index_add_)."""
import json, os, sys
sys.path.insert(0, ".")
import torch
import torch.distributed as dist
from kubedl_b200.worker.ddp_trainer import init_distributed
from kubedl_b200.models.din import DINModel, synthetic_batch, D

rank, world, _ = init_distributed("nccl")
dev = torch.device("cuda", torch.cuda.current_device())
V, L_, B = int(os.environ.get("VOCAB", 10_000_000)), 62, int(os.environ.get("BATCH", 4096))
STEPS = 20


def timed(fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(STEPS): fn()
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / STEPS], device=dev)
    if world > 1: dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return ms.item()


batches = [synthetic_batch(B, L_, V, dev, seed=1000 * rank + i) for i in range(4)]
m = DINModel(vocab=V, hist_len=L_, emb_lr=20.0)
it = [0]
def ours():
    m.step(*batches[it[0] % 4]); it[0] += 1
t_ours = timed(ours)
# exchange-only timing of the two fused kernels
x = torch.empty(B, 4 * D, dtype=torch.bfloat16, device=dev)
def ours_exchange():
    u, c_, h, _ = batches[it[0] % 4]; it[0] += 1
    m.forward(u, c_, h)
    # backward exchange with a synthetic upstream gradient
    user, cand, hist, xs, hs, aw, h1, h2 = m.saved; m.saved = None
    from kubedl_b200.ops import lib as L
    L.check(m.lib.kdl_din_scatter_update(m._shard_tab, m.world, m.rows_per_rank, m.interleaved, user.data_ptr(), cand.data_ptr(),
                                         hist.data_ptr(), B, L_, D, xs.data_ptr(), x.data_ptr(), hs.data_ptr(),
                                         aw.data_ptr(), 0.0, L.stream_ptr()))
    if m.ar is not None: m.ar(m.arena.g16_off, 8, torch.bfloat16, channel=6)
x.normal_()
t_ours_x = timed(ours_exchange)

# ---- stock baseline: all_to_all id / vector exchange around a local table shard
rows_per_rank = m.rows_per_rank
table = m.shard.clone()
mlp = [torch.nn.Linear(4 * D, 512), torch.nn.Linear(512, 256), torch.nn.Linear(256, 1)]
mlp = torch.nn.Sequential(mlp[0], torch.nn.ReLU(), mlp[1], torch.nn.ReLU(), mlp[2]).to(dev).bfloat16()
opt = torch.optim.SGD(mlp.parameters(), lr=0.05)

def exchange_ids(ids):
    owner = m.owner_of(ids)              # same row placement as the fused path
    order = torch.argsort(owner, stable=True)
    sorted_ids = ids[order]
    send_counts = torch.bincount(owner, minlength=world)
    recv_counts = torch.empty_like(send_counts)
    if world > 1: dist.all_to_all_single(recv_counts, send_counts)
    else: recv_counts.copy_(send_counts)
    sc, rc = send_counts.tolist(), recv_counts.tolist()
    recv_ids = torch.empty(sum(rc), dtype=ids.dtype, device=dev)
    if world > 1: dist.all_to_all_single(recv_ids, sorted_ids, rc, sc)
    else: recv_ids.copy_(sorted_ids)
    return order, sc, rc, recv_ids

def stock():
    user, cand, hist, label = batches[it[0] % 4]; it[0] += 1
    mask = hist >= 0
    ids = torch.cat([user, cand, hist[mask]])
    order, sc, rc, recv_ids = exchange_ids(ids)
    local = m.local_of(recv_ids)
    rows = table.index_select(0, local)                                   # owner-side gather
    back = torch.empty(ids.numel(), D, dtype=rows.dtype, device=dev)
    if world > 1: dist.all_to_all_single(back, rows, sc, rc)
    else: back.copy_(rows)
    vec = torch.empty_like(back); vec[order] = back
    vec.requires_grad_(True)
    u, c_ = vec[:B], vec[B:2 * B]
    h = torch.zeros(B, L_, D, device=dev, dtype=vec.dtype).masked_scatter(mask[..., None].expand(-1, -1, D), vec[2 * B:])
    s = (h * c_[:, None]).sum(-1) / D ** 0.5
    a = torch.softmax(s.masked_fill(~mask, float("-inf")), 1)
    p = (a[..., None] * h).sum(1)
    logit = mlp(torch.cat([u, c_, p, p * c_], 1).bfloat16())[:, 0].float()
    loss = torch.nn.functional.binary_cross_entropy_with_logits(logit, label)
    opt.zero_grad(); loss.backward()
    g = vec.grad[order].contiguous()                                      # back to the owners
    grecv = torch.empty(sum(rc), D, dtype=g.dtype, device=dev)
    if world > 1: dist.all_to_all_single(grecv, g, rc, sc)
    else: grecv.copy_(g)
    table.index_add_(0, local, grecv, alpha=-20.0)
    if world > 1:
        for q in mlp.parameters(): dist.all_reduce(q.grad)
    opt.step()
t_stock = timed(stock)
lookups = B * (2 + L_)
if rank == 0:
    print(json.dumps({"world": world, "vocab": V, "batch_per_gpu": B, "hist": L_,
                      "ours_ms_per_step": t_ours, "ours_samples_per_s": B * world / (t_ours / 1e3),
                      "ours_exchange_only_ms": t_ours_x,
                      "stock_ms_per_step": t_stock, "stock_samples_per_s": B * world / (t_stock / 1e3),
                      "speedup": t_stock / t_ours,
                      "vector_bytes_per_gpu_per_step": lookups * D * 4 * 2}), flush=True)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
