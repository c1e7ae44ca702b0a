"""PyTorchJob example worker (BASELINE.json config #1): MNIST-shaped DDP training over gloo on
CPU, wired purely by the env the operator injects (MASTER_ADDR / MASTER_PORT / WORLD_SIZE /
RANK) — the local stand-in for the `kubedl/pytorch-dist-example` image of reference
example/pytorch/pytorch_job_mnist_mpi.yaml:15.  Synthetic data (no dataset download)."""
from __future__ import annotations

import argparse
import os
import sys
import time


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=int(os.environ.get("KUBEDL_MNIST_STEPS", "20")))
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--original-command", default="")
    ap.add_argument("--fail-with", type=int, default=int(os.environ.get("KUBEDL_FAIL_WITH", "0")))
    ap.add_argument("--checkpoint", default=os.environ.get("KUBEDL_CHECKPOINT", ""))
    args, _ = ap.parse_known_args(argv)
    import torch
    import torch.distributed as dist
    from .rendezvous import init_process_group
    t0 = time.time()
    rank, world = init_process_group("gloo", "pytorch")
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(784, 128), torch.nn.ReLU(),
                                torch.nn.Linear(128, 10))
    start_step = 0
    if args.checkpoint and os.path.exists(args.checkpoint):
        sd = torch.load(args.checkpoint, weights_only=False)
        model.load_state_dict(sd["model"])
        start_step = sd["step"]
    # Data parallelism by hand: one flat gradient all-reduce per step.  torch's DistributedDataParallel
    # would do the same for this model, but constructing it imports torch._dynamo (5 s on this
    # image), which is most of the job's wall time.
    params = list(model.parameters())
    if world > 1:
        flat = torch.cat([p.detach().reshape(-1) for p in params])
        dist.broadcast(flat, src=0)                    # identical start, whatever the seeds did
        o = 0
        with torch.no_grad():
            for p in params:
                p.copy_(flat[o:o + p.numel()].view_as(p))
                o += p.numel()
    ddp = model
    lr = 0.05        # plain SGD by hand: torch.optim pulls in torch._dynamo as well
    g = torch.Generator().manual_seed(100 + rank)
    loss = None
    for step in range(start_step, args.steps):
        x = torch.randn(args.batch, 1, 28, 28, generator=g)
        y = (x.flatten(1)[:, :10]).argmax(1)  # learnable synthetic labels
        loss = torch.nn.functional.cross_entropy(ddp(x), y)
        for p in params:
            p.grad = None
        loss.backward()
        if world > 1:
            g_flat = torch.cat([p.grad.reshape(-1) for p in params])
            dist.all_reduce(g_flat)
            g_flat /= world
            o = 0
            for p in params:
                p.grad.copy_(g_flat[o:o + p.numel()].view_as(p))
                o += p.numel()
        with torch.no_grad():
            for p in params:
                p.add_(p.grad, alpha=-lr)
        once = os.environ.get("KUBEDL_FAIL_ONCE", "")      # marker file: only the first incarnation fails
        if args.fail_with and step == 2 and rank == world - 1 and not os.environ.get("KUBEDL_RESTARTED") \
                and not (once and os.path.exists(once)):
            if once:
                open(once, "w").close()
            print(f"rank {rank}: injected failure with exit code {args.fail_with}", flush=True)
            os._exit(args.fail_with)
        if args.checkpoint and rank == 0 and step % 5 == 4:
            torch.save({"model": model.state_dict(), "step": step + 1}, args.checkpoint)
    print(f"rank {rank}/{world}: {args.steps - start_step} steps, loss {float(loss.detach()) if loss is not None else float('nan'):.4f}, "
          f"time-to-done {time.time() - t0:.2f}s", flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    from kubedl_b200.worker.exitcodes import run
    sys.exit(run(main))
