"""Gather (N4) and scatter-update (N5) kernels timed separately at the BASELINE config
(10M x 64 fp32 table over the ranks, B = 4096 per GPU, 62-item histories)."""
import json, os, sys
sys.path.insert(0, ".")
import torch
import torch.distributed as dist
from kubedl_b200.worker.ddp_trainer import init_distributed
from kubedl_b200.models.din import DINModel, synthetic_batch, D
from kubedl_b200.ops import lib as L

rank, world, _ = init_distributed("nccl")
dev = torch.device("cuda", torch.cuda.current_device())
V, L_, B = 10_000_000, 62, 4096
m = DINModel(vocab=V, hist_len=L_, emb_lr=0.0)
batches = [synthetic_batch(B, L_, V, dev, seed=1000 * rank + i) for i in range(4)]
dx = torch.randn(B, 4 * D, device=dev).bfloat16()


def timed(fn, steps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps): fn()
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    if world > 1: dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return ms.item()

it = [0]
def gather():
    u, c, h, _ = batches[it[0] % 4]; it[0] += 1
    m.forward(u, c, h)          # gather kernel + 3 small GEMMs
def gather_only():
    u, c, h, _ = batches[it[0] % 4]; it[0] += 1
    x = torch.empty(B, 4 * D, dtype=torch.bfloat16, device=dev)
    hs = torch.empty(B, L_, D, dtype=torch.bfloat16, device=dev)
    aw = torch.empty(B, L_, dtype=torch.float32, device=dev)
    L.check(m.lib.kdl_din_gather_attend(m._shard_tab, m.world, m.rows_per_rank, m.interleaved, u.data_ptr(), c.data_ptr(), h.data_ptr(),
                                        B, L_, D, x.data_ptr(), hs.data_ptr(), aw.data_ptr(), L.stream_ptr()))
    m._last = (u, c, h, x, hs, aw)
def scatter_only():
    u, c, h, x, hs, aw = m._last
    L.check(m.lib.kdl_din_scatter_update(m._shard_tab, m.world, m.rows_per_rank, m.interleaved, u.data_ptr(), c.data_ptr(), h.data_ptr(),
                                         B, L_, D, x.data_ptr(), dx.data_ptr(), hs.data_ptr(), aw.data_ptr(), 0.0,
                                         L.stream_ptr()))
res = {"world": world, "gather_ms": timed(gather_only)}
gather_only()
res["scatter_ms"] = timed(scatter_only)
# bytes that really cross NVLink per kernel: VALID rows (histories are variable-length, -1 = padding) whose owner
# is another rank, 256 bytes each, averaged over the batches used; the link roofline is 770 GB/s per direction
def remote_rows(b):
    u, c, h, _ = b
    ids = torch.cat([u, c, h.flatten()])
    ids = ids[ids >= 0]
    return int((m.owner_of(ids) != rank).sum()), int(ids.numel())
rr = [remote_rows(b) for b in batches]
remote_mb = sum(r for r, _ in rr) / len(rr) * D * 4 / 1e6
t = torch.tensor([remote_mb], device=dev)
if world > 1: dist.all_reduce(t, op=dist.ReduceOp.MAX)
res["rows_per_kernel"] = sum(n for _, n in rr) / len(rr)
res["remote_mb_per_kernel"] = float(t.item())
res["remote_mb_if_all_slots_were_valid"] = B * (L_ + 2) * D * 4 * (world - 1) / world / 1e6
for k in ("gather", "scatter"):
    gbs = res["remote_mb_per_kernel"] / res[k + "_ms"] if world > 1 else 0.0
    res[k + "_remote_gb_s"] = gbs
    res[k + "_frac_of_770"] = gbs / 770.0
res["gather_impl"] = os.environ.get("KDL_DIN_GATHER", "tma")
res["sharding"] = m.sharding
# egress of the busiest OWNER: rows every other rank pulls from one shard
own = torch.zeros(world, device=dev)
for b in batches:
    u, c, h, _ = b
    ids = torch.cat([u, c, h.flatten()]); ids = ids[ids >= 0]
    o = m.owner_of(ids)
    own += torch.bincount(o[o != rank], minlength=world).float() / len(batches)
if world > 1: dist.all_reduce(own)
res["busiest_owner_egress_mb"] = float(own.max()) * D * 4 / 1e6
res["gather_busiest_port_gb_s"] = res["busiest_owner_egress_mb"] / res["gather_ms"]
if rank == 0:
    print(json.dumps(res), flush=True)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
