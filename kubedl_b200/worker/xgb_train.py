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


def model_file(args) -> str:
    """Where `--model_path` lives on this box.  The reference image writes to OSS or the pod's
    disk (`--model_storage_type=oss|local`, example/xgboost/xgboostjob_v1alpha1_iris_train.yaml:26-28);
    there is no object store here, so both land under $KUBEDL_MODEL_DIR (a shared volume in a
    real deployment) and the log says so."""
    if not args.model_path:
        return ""
    root = os.environ.get("KUBEDL_MODEL_DIR", "/tmp/kubedl-b200/models")
    rel = args.model_path.lstrip("/") if args.model_storage_type != "local" or not os.path.isabs(args.model_path) \
        else args.model_path
    path = rel if os.path.isabs(rel) else os.path.join(root, rel)
    return path if path.endswith(".pt") else os.path.join(path, "model.pt")


def synthetic_shard(rows: int, features: int, rank: int, dev):
    import torch
    g = torch.Generator(device=dev).manual_seed(7 + rank)
    X = torch.randn(rows, features, device=dev, generator=g)
    w = torch.linspace(-1, 1, features, device=dev)
    y = (torch.rand(rows, device=dev, generator=g) < torch.sigmoid(X @ w + X[:, 0] * X[:, 1])).float()
    return X, y


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--job_type", default="Train", help="Train | Predict")
    ap.add_argument("--n_estimators", type=int, default=10)
    ap.add_argument("--learning_rate", type=float, default=0.1)
    ap.add_argument("--max_depth", type=int, default=int(os.environ.get("KUBEDL_XGB_DEPTH", "6")))
    ap.add_argument("--rows", type=int, default=int(os.environ.get("KUBEDL_XGB_ROWS", "200000")), help="rows per rank")
    ap.add_argument("--features", type=int, default=28)
    ap.add_argument("--comm", default=os.environ.get("KUBEDL_XGB_COMM", "fused"))
    ap.add_argument("--model_path", default=os.environ.get("KUBEDL_XGB_MODEL_PATH", ""))
    ap.add_argument("--model_storage_type", default="local")
    ap.add_argument("--original-command", default="")
    args, _ = ap.parse_known_args(argv)
    import torch
    import torch.distributed as dist
    from ..models import gbdt
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
    X, y = synthetic_shard(args.rows, args.features, rank, dev)
    path = model_file(args)
    t0 = time.time()
    out = {"rank": rank, "world": world, "job_type": args.job_type, "device": str(dev)}
    if args.job_type.lower() == "predict":
        if not path or not os.path.exists(path):
            print(f"Predict needs the model a Train job wrote; nothing at {path or '--model_path'}", flush=True)
            return 2
        model = gbdt.load_model(path)
        margin = gbdt.predict_margin(model, X)
        stats = torch.stack([torch.nn.functional.binary_cross_entropy_with_logits(margin, y, reduction="sum"),
                             ((margin > 0).float() == y).float().sum(), torch.tensor(float(len(y)), device=dev)])
        if world > 1:
            dist.all_reduce(stats)
        out.update(trees=len(model["trees"]), logloss=float(stats[0] / stats[2]), accuracy=float(stats[1] / stats[2]),
                   rows=int(stats[2]), model=path)
    else:
        if use_cuda:
            tr = gbdt.GBDTTrainer(X, y, max_depth=args.max_depth, learning_rate=args.learning_rate, comm=args.comm)
            for _ in range(args.n_estimators):
                tr.grow_tree()
            loss, cuts, trees = float(tr.logloss()), tr.cuts, tr.trees
        else:
            # no GPU: the same algorithm on the PyTorch oracle, histograms summed over the ranks
            # with gloo (small trees: this path exists so the example runs anywhere)
            depth, n_trees = min(args.max_depth, 4), min(args.n_estimators, int(os.environ.get("KUBEDL_XGB_CPU_TREES", "3")))
            cuts = gbdt.compute_cuts(X)
            bins = gbdt.quantize(X, cuts)
            pred = torch.zeros(args.rows)
            reduce = (lambda t: dist.all_reduce(t)) if world > 1 else None
            trees = []
            for _ in range(n_trees):
                p = torch.sigmoid(pred)
                gh = torch.stack([p - y, (p * (1 - p)).clamp_min(1e-16)], 1)
                splits, leaf, node = gbdt.reference_tree_torch(bins, gh, args.features, depth, 1.0, args.learning_rate,
                                                               allreduce=reduce)
                trees.append(gbdt.tree_from_reference(splits, leaf))
                pred += leaf[node].float()
            loss = float(torch.nn.functional.binary_cross_entropy_with_logits(pred, y))
        out.update(trees=len(trees), logloss=loss)
        if path and rank == 0:
            gbdt.save_model(path, gbdt.export_model(cuts, trees, learning_rate=args.learning_rate, features=args.features))
            note = "" if args.model_storage_type == "local" else f" (storage type {args.model_storage_type!r} has no backend here: kept on local disk)"
            print(f"model with {len(trees)} trees written to {path}{note}", flush=True)
            out["model"] = path
    out["seconds"] = time.time() - t0
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    from kubedl_b200.worker.exitcodes import run
    sys.exit(run(main))
