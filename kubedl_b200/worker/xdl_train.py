"""XDLJob worker (BASELINE.json config #5): DIN-style sparse CTR training with the row-sharded
embedding table exchanged over NVLink by the fused gather / scatter kernels.  Roles follow the
XDLJob contract (TASK_NAME / TASK_INDEX, reference controllers/xdl/xdljob_controller.go:191-217):
`worker` replicas train; `ps` and `scheduler` replicas have nothing to do here — the "parameter
server" is the symmetric heap itself — so they idle until the workers are done (scheduler) or
forever (ps; cleaned up by CleanPodPolicy Running).  ZooKeeper discovery is replaced by a
file-based rendezvous under the ZK_ADDR path name.  On a CPU-only box a tiny dense stand-in runs."""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import sys
import time


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=int(os.environ.get("KUBEDL_XDL_STEPS", "20")))
    ap.add_argument("--batch", type=int, default=int(os.environ.get("KUBEDL_XDL_BATCH", "4096")))
    ap.add_argument("--vocab", type=int, default=int(os.environ.get("KUBEDL_XDL_VOCAB", "10000000")))
    ap.add_argument("--hist", type=int, default=62)
    ap.add_argument("--workers", type=int, default=int(os.environ.get("KUBEDL_XDL_WORKERS", "0")))
    ap.add_argument("--original-command", default="")
    args, _ = ap.parse_known_args(argv)
    from .rendezvous import xdl_identity
    task, index, zk = xdl_identity()
    if task in ("ps", "scheduler", "extendrole"):
        print(f"{task}:{index} idle (embedding shards live in the workers' symmetric heap)", flush=True)
        if task == "ps":
            time.sleep(float(os.environ.get("KUBEDL_XDL_PS_SECONDS", "3600")))
        else:
            time.sleep(float(os.environ.get("KUBEDL_XDL_SCHED_SECONDS", "0.5")))
        return 0
    import torch
    import torch.distributed as dist
    world = args.workers or int(os.environ.get("WORLD_SIZE", "1"))
    use_cuda = torch.cuda.is_available()
    if world > 1:
        # rendezvous port derived from the job-unique ZK path (ZK_ADDR carries /<job uid>)
        port = 20000 + int(hashlib.sha1(zk.encode()).hexdigest(), 16) % 20000
        if use_cuda:
            torch.cuda.set_device(int(os.environ.get("KUBEDL_GPU_INDEX", str(index))) % torch.cuda.device_count())
        dist.init_process_group("nccl" if use_cuda else "gloo", init_method=f"tcp://127.0.0.1:{port}", rank=index, world_size=world)
    t0 = time.time()
    if use_cuda:
        from ..models.din import DINModel, synthetic_batch
        dev = torch.device("cuda", torch.cuda.current_device())
        m = DINModel(vocab=args.vocab, hist_len=args.hist, emb_lr=20.0)
        loss = None
        for it in range(args.steps):
            loss = m.step(*synthetic_batch(args.batch, args.hist, args.vocab, dev, seed=it * 131 + index))
        loss = float(loss)
    else:
        emb = torch.nn.Embedding(1000, 16)
        lin = torch.nn.Linear(16, 1)
        params = list(emb.parameters()) + list(lin.parameters())
        for it in range(args.steps):
            ids = torch.randint(0, 1000, (256,))
            loss = torch.nn.functional.binary_cross_entropy_with_logits(lin(emb(ids))[:, 0], (ids % 2).float())
            for p in params:
                p.grad = None
            loss.backward()
            with torch.no_grad():          # SGD by hand: torch.optim imports torch._dynamo (5 s)
                for p in params:
                    p.add_(p.grad, alpha=-0.5)
        loss = float(loss.detach())
    print(json.dumps({"task": task, "index": index, "world": world, "steps": args.steps, "loss": loss,
                      "seconds": time.time() - t0, "zk_addr": zk}), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    from kubedl_b200.worker.exitcodes import run
    sys.exit(run(main))
