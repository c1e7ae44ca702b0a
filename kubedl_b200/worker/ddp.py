"""PyTorchJob GPU worker (BASELINE.json config #2): ResNet-50 DDP bf16 on synthetic 224² images
through kubedl_b200.worker.ddp_trainer.DDPTrainer, one process per GPU, wired by MASTER_ADDR /
MASTER_PORT / WORLD_SIZE / RANK.  Prints samples/s and time-to-first-step."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=int(os.environ.get("KUBEDL_DDP_STEPS", "20")))
    ap.add_argument("--batch", type=int, default=int(os.environ.get("KUBEDL_DDP_BATCH", "256")))
    ap.add_argument("--image-size", type=int, default=224)
    ap.add_argument("--comm", default=os.environ.get("KUBEDL_DDP_COMM", "fused"))
    ap.add_argument("--checkpoint", default=os.environ.get("KUBEDL_CHECKPOINT", ""))
    ap.add_argument("--original-command", default="")
    args, _ = ap.parse_known_args(argv)
    t_start = time.time()
    import torch
    from .ddp_trainer import DDPTrainer
    t_import = time.time() - t_start
    tr = DDPTrainer(batch_per_gpu=args.batch, comm=args.comm, image_size=args.image_size)
    t_init = time.time() - t_start
    if args.checkpoint:
        tr.load(args.checkpoint)
    g = torch.Generator().manual_seed(tr.rank)
    img = torch.randint(0, 256, (args.batch, args.image_size, args.image_size, 3), dtype=torch.uint8, generator=g).pin_memory()
    lab = torch.randint(0, 1000, (args.batch,), dtype=torch.int64, generator=g).pin_memory()
    tr.train_step(img, lab, sync=True)
    t_first = time.time() - t_start
    # what the user waits for: from the node agent starting this pod (fork / exec / adoption of a waiting worker)
    pod_start = float(os.environ.get("KUBEDL_POD_START_TS", "0") or 0)
    t_pod = (time.time() - pod_start) if pod_start else None
    tr.warmup(2)
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(args.steps):
        tr.train_step(img, lab)
    loss = tr.last_loss()
    dt = time.time() - t0
    if args.checkpoint:
        tr.save(args.checkpoint)
    print(json.dumps({"rank": tr.rank, "world": tr.world, "samples_per_s_per_gpu": args.batch * args.steps / dt,
                      "time_to_first_step_s": t_first,
                      "pod_start_to_first_step_s": t_pod, "loss": loss, "comm": tr.comm,
                      "startup_s": {"import": round(t_import, 3), **tr.startup,
                                    "first_step": round(t_first - t_init, 3)}}), flush=True)
    if tr.world > 1:
        import torch.distributed as dist
        dist.barrier(); dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    from kubedl_b200.worker.exitcodes import run
    sys.exit(run(main))
