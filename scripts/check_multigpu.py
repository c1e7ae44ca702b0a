"""Multi-GPU check of the symmetric heap + fused all-reduce (run under torchrun)."""
import json, os, sys, time
sys.path.insert(0, ".")
import torch
import torch.distributed as dist
from kubedl_b200.worker.ddp_trainer import init_distributed
from kubedl_b200.parallel.symm import SymmetricHeap
from kubedl_b200.parallel.collectives import FusedAllReduce

rank, world, _ = init_distributed("nccl")
dev = torch.device("cuda", torch.cuda.current_device())
mode = os.environ.get("KUBEDL_SYMM_MODE", "auto")
heap = SymmetricHeap(512 << 20, mode=mode)
if rank == 0:
    print(json.dumps({"transport": heap.transport, "nvls": bool(heap.mc_ptr), "world": world}), flush=True)
res = {}
for use_nvls in ([True, False] if heap.mc_ptr else [False]):
    ar = FusedAllReduce(heap, use_nvls=use_nvls)
    tag = "nvls" if use_nvls else "p2p"
    for n in (8 * 1024, 4 << 20, 32 << 20):     # elements (bf16): 16 KB, 8 MB, 64 MB
        off, t = heap.alloc_tensor((n,), torch.bfloat16)
        torch.manual_seed(100 + rank)
        src = torch.randn(n, device=dev).bfloat16()
        t.copy_(src)
        ref = src.float().clone()
        dist.all_reduce(ref)
        ref /= world
        torch.cuda.synchronize(); dist.barrier()
        ar(off, n)
        torch.cuda.synchronize()
        err = (t.float() - ref).abs().max().item()
        gathered = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(gathered, t)
        same = all(torch.equal(gathered[0], g) for g in gathered)
        # timing: fused vs NCCL all_reduce + scale
        def bench(fn, iters=20):
            for _ in range(3): fn()
            torch.cuda.synchronize(); dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters): fn()
            e1.record(); torch.cuda.synchronize()
            ms = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            return ms.item()
        t_f = bench(lambda: ar(off, n))
        buf = src.clone()
        t_n = bench(lambda: (dist.all_reduce(buf), buf.mul_(1.0 / world)))
        busbw = 2 * (world - 1) / world * n * 2 / (t_f * 1e-3) / 1e9
        res[f"{tag}_{n}"] = {"max_err": err, "bitwise_same_on_all_ranks": same, "fused_ms": t_f, "nccl_ms": t_n,
                             "fused_busbw_GBs": busbw}
        assert err < 5e-2 and same, (tag, n, err, same)
    # fp32 -> bf16 fused cast path (P2P only)
    if not use_nvls:
        n = 1 << 20
        off_in, tin = heap.alloc_tensor((n,), torch.float32)
        off_out, tout = heap.alloc_tensor((n,), torch.bfloat16)
        src = torch.randn(n, device=dev)
        tin.copy_(src)
        ref = src.clone(); dist.all_reduce(ref); ref /= world
        torch.cuda.synchronize(); dist.barrier()
        ar(off_in, n, torch.float32, out_off=off_out, out_dtype=torch.bfloat16)
        torch.cuda.synchronize()
        e = (tout.float() - ref).abs().max().item()
        assert e < 5e-2, e
        res["f32_to_bf16_err"] = e
assert heap.status_word() == 0
if rank == 0:
    print(json.dumps(res), flush=True)
dist.barrier()
dist.destroy_process_group()
