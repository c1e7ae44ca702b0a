"""Distributed histogram GBDT (BASELINE.json config #3: "XGBoostJob distributed hist training ...
synthetic HIGGS-shape 11M×28, fused histogram allreduce").

Data parallel over rows: every rank holds a shard, quantises it to 256 bins per feature with
cuts agreed on by all ranks, and grows depth-wise trees level by level.  Per level:

  rows grouped by node -> global row counts (fused all-reduce) -> build the SMALLER child of every
  parent with `kdl_gbdt_hist_allreduce` (shared-memory histograms + in-kernel one-shot cross-GPU
  reduction over NVLink) -> sibling = parent − built (subtraction trick) -> `kdl_gbdt_split_eval`
  -> `kdl_gbdt_apply_split`.

Every rank sees bitwise-identical global histograms (fixed reduction order), hence grows the same
tree without exchanging split decisions.  Stock equivalent: XGBoost `tree_method=hist` on GPU with
Rabit/NCCL all-reduce of histograms, wired by MASTER_ADDR/RANK from the XGBoostJob controller
(reference controllers/xgboost/pod.go:106-152).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional

import torch

from ..ops import lib as L

L.register_signatures({
    "kdl_gbdt_grad_logistic": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "kdl_gbdt_hist_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "kdl_gbdt_hist_allreduce": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_void_p, C.c_int, C.c_int,
                                          C.c_int, C.c_uint, C.c_void_p, C.c_ulonglong, C.c_float, C.c_void_p,
                                          C.c_void_p]),
    "kdl_gbdt_split_eval": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float,
                                      C.c_void_p, C.c_void_p]),
    "kdl_gbdt_apply_split": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p]),
    "kdl_gbdt_update_pred": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "kdl_gbdt_choose_build": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "kdl_gbdt_assemble_hist": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "kdl_gbdt_leaf_values": (C.c_int, [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]),
    "kdl_gbdt_partition_workspace_bytes": (C.c_size_t, [C.c_int]),
    "kdl_gbdt_partition": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_size_t, C.c_void_p]),
})

FSTRIDE = 32
NBINS = 256


@dataclass
class Tree:
    depth: int
    splits: List[torch.Tensor] = field(default_factory=list)   # per level: [2^d, 8] (feat, thr, gain, GL, HL, G, H, valid)
    leaf_value: Optional[torch.Tensor] = None                  # [2^depth]


def compute_cuts(X: torch.Tensor, group=None, sample: int = 200_000) -> torch.Tensor:
    """255 cut points per feature from a sample gathered over all ranks (identical everywhere)."""
    import torch.distributed as dist
    n = X.shape[0]
    g = torch.Generator(device=X.device).manual_seed(1234)
    idx = torch.randint(0, n, (min(sample, n),), device=X.device, generator=g)
    s = X[idx]
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        parts = [torch.empty_like(s) for _ in range(dist.get_world_size(group))]
        dist.all_gather(parts, s.contiguous(), group=group)
        s = torch.cat(parts, 0)
    q = torch.linspace(0, 1, NBINS + 1, device=X.device, dtype=torch.float32)[1:-1]
    return torch.quantile(s.float(), q, dim=0).t().contiguous()        # [F, 255]


def quantize(X: torch.Tensor, cuts: torch.Tensor) -> torch.Tensor:
    n, F = X.shape
    bins = torch.zeros(n, FSTRIDE, dtype=torch.uint8, device=X.device)
    for f in range(F):
        bins[:, f] = torch.bucketize(X[:, f].contiguous(), cuts[f].contiguous(), right=False).to(torch.uint8)
    return bins


class GBDTTrainer:
    def __init__(self, X: torch.Tensor, y: torch.Tensor, max_depth: int = 6, learning_rate: float = 0.1,
                 reg_lambda: float = 1.0, gamma: float = 0.0, min_child_weight: float = 1.0, heap=None,
                 cuts: Optional[torch.Tensor] = None, group=None, comm: str = "fused", use_graph: bool = True):
        import torch.distributed as dist
        self.lib = L.lib()
        self.dev = X.device
        self.n, self.F = X.shape
        assert self.F <= 32
        self.depth, self.lr, self.lam, self.gamma, self.mcw = max_depth, learning_rate, reg_lambda, gamma, min_child_weight
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.heap = heap
        self.comm = comm if self.world > 1 else "none"
        self.group = group
        import os as _os
        self.use_nvls = _os.environ.get("KUBEDL_NVLS", "auto") != "0"
        if self.world > 1 and heap is None and comm == "fused":
            from ..parallel.symm import SymmetricHeap
            self.heap = heap = SymmetricHeap(64 << 20, group=group)
        self.cuts = cuts if cuts is not None else compute_cuts(X, group)
        self.bins = quantize(X, self.cuts)
        self.y = y.float().contiguous()
        self.pred = torch.zeros(self.n, dtype=torch.float32, device=self.dev)
        self.gh = torch.empty(self.n, 2, dtype=torch.float32, device=self.dev)
        self.node = torch.zeros(self.n, dtype=torch.int32, device=self.dev)
        self.max_nodes = 1 << max(max_depth - 1, 0)            # most nodes ever built in one level
        hist_floats = self.max_nodes * self.F * NBINS * 2
        if heap is not None:
            # two level-parity buffers: a buffer is rewritten only after every peer has passed the
            # next level's in-kernel barrier, i.e. finished reading it
            self.local_off = [heap.alloc(hist_floats * 4) for _ in range(2)]
            self.local = [heap.tensor(o, (hist_floats,), torch.float32) for o in self.local_off]
            self.flags_off = heap.alloc(8 * self.max_nodes * 4 * 2)
            self.cnt_off, self.cnt = heap.alloc_tensor((max(256, 1 << max_depth),), torch.float32)
            from ..parallel.collectives import FusedAllReduce
            self.ar = FusedAllReduce(heap, use_nvls=False)
            self._peer_tabs = [heap.ptr_table(o) for o in self.local_off]
            self._flag_tab = heap.ptr_table(self.flags_off)
        else:
            self.local = [torch.zeros(hist_floats, dtype=torch.float32, device=self.dev) for _ in range(2)]
        self.out = torch.empty(hist_floats, dtype=torch.float32, device=self.dev)
        self.node_done = torch.zeros(self.max_nodes, dtype=torch.int32, device=self.dev)
        self.epoch = 0
        self.level_parity = 0
        self.trees: List[Tree] = []
        # {epoch, finished nodes} of the in-kernel cross-GPU barrier live on the device, so a whole
        # boosting round can be captured into a CUDA graph and replayed
        self.epoch_state = torch.zeros(2, dtype=torch.int32, device=self.dev)
        self.use_graph = use_graph and self.comm != "nccl"
        self._graph = None
        self.gh_abs_max = 1.0    # logistic loss: |g| <= 1, h <= 1/4 (bound of the fixed-point histogram scale)

    # ------------------------------------------------------------------ one tree
    def _global_counts(self, counts: torch.Tensor) -> torch.Tensor:
        if self.world == 1:
            return counts
        if self.comm == "nccl":
            import torch.distributed as dist
            c2 = counts.clone()
            dist.all_reduce(c2, group=self.group)
            return c2
        n = self.cnt.numel()
        self.cnt.zero_()
        self.cnt[:counts.numel()] = counts.float()
        self.ar(self.cnt_off, n, torch.float32, scale=1.0, channel=7)
        return self.cnt[:counts.numel()]      # consumed (stream-ordered) before the next level zeroes it

    def _partition(self, nodes: int):
        """Rows grouped by node (one counting-sort pass, csrc/gbdt.cu) -> (ridx, seg, counts)."""
        lib = self.lib
        if not hasattr(self, "_ridx"):
            self._ridx = torch.empty(self.n, dtype=torch.int32, device=self.dev)
            nb = lib.kdl_gbdt_partition_workspace_bytes(1 << max(self.depth - 1, 0))
            self._part_ws = torch.empty(nb, dtype=torch.uint8, device=self.dev)
        seg = torch.empty(nodes + 1, dtype=torch.int32, device=self.dev)
        counts = torch.empty(nodes, dtype=torch.int32, device=self.dev)
        L.check(lib.kdl_gbdt_partition(self.node.data_ptr(), self.n, nodes, self._ridx.data_ptr(), seg.data_ptr(),
                                       counts.data_ptr(), self._part_ws.data_ptr(), self._part_ws.numel(),
                                       L.stream_ptr()), "kdl_gbdt_partition")
        L.count(4)
        return self._ridx, seg, counts

    def _build_hist(self, ridx, seg, build, nbuild) -> torch.Tensor:
        par = self.level_parity
        self.level_parity ^= 1
        hist_floats = nbuild * self.F * NBINS * 2
        self.local[par][:hist_floats].zero_()
        self.epoch += 1
        fused = self.heap is not None and self.comm == "fused"
        rc = self.lib.kdl_gbdt_hist_allreduce(
            self.bins.data_ptr(), FSTRIDE, self.F, self.gh.data_ptr(), ridx.data_ptr(), seg.data_ptr(),
            build.data_ptr(), nbuild, self.n, self.local[par].data_ptr(), self.out.data_ptr(),
            self.node_done.data_ptr(), self._peer_tabs[par] if fused else None,
            self._flag_tab if fused else None,
            (self.heap.mc(self.local_off[par]) if (fused and self.use_nvls) else None),
            self.world if fused else 1, self.rank if fused else 0,
            self.max_nodes, self.epoch, None, int(30e9), self.gh_abs_max,
            self.epoch_state.data_ptr() if fused else None, L.stream_ptr())
        L.check(rc, "kdl_gbdt_hist_allreduce")
        L.count(1)
        out = self.out[:hist_floats]
        if self.comm == "nccl":   # baseline: separate NCCL all-reduce of the finished histograms
            import torch.distributed as dist
            dist.all_reduce(out, group=self.group)
        return out.view(nbuild, self.F, NBINS, 2)

    def grow_tree(self) -> Tree:
        """One boosting round.  The launch sequence of a round is static (≈150 launches, no host
        synchronisation), so after one eager round it is captured into a CUDA graph and replayed:
        at 8 GPUs a round is otherwise bound by launch latency, not by the kernels."""
        if not self.use_graph:
            tree = self._grow_tree_eager()
        elif self._graph is None and not getattr(self, "_warm", False):
            self._warm = True
            tree = self._grow_tree_eager()
        else:
            if self._graph is None:
                L.lib()                      # everything is loaded before capture
                torch.cuda.synchronize()
                self._graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self._graph):
                    self._graph_tree = self._grow_tree_eager()
                self.trees.pop()             # capture does not execute: that tree never existed
            self._graph.replay()
            L.count(self._graph_launches)
            src = self._graph_tree
            tree = Tree(src.depth, [t.clone() for t in src.splits], src.leaf_value.clone())
            self.trees.append(tree)
        return tree

    def _grow_tree_eager(self) -> Tree:
        n0 = L.launch_count()
        tree = self._grow_tree_impl()
        self._graph_launches = L.launch_count() - n0
        return tree

    def _grow_tree_impl(self) -> Tree:
        lib = self.lib
        L.check(lib.kdl_gbdt_grad_logistic(self.pred.data_ptr(), self.y.data_ptr(), self.gh.data_ptr(), self.n,
                                           L.stream_ptr()), "kdl_gbdt_grad_logistic")
        L.count(1)
        self.node.zero_()
        tree = Tree(self.depth)
        parent_hist = None
        if self.world > 1 and self.comm == "fused":
            # cross-rank barrier: nobody zeroes a histogram buffer a peer may still be reading from the
            # previous round's last level (inside a round the row-count all-reduce of every level does it)
            self._global_counts(torch.zeros(1, dtype=torch.int64, device=self.dev))
        self.level_parity = 0
        for d in range(self.depth):
            nodes = 1 << d
            ridx, seg, counts = self._partition(nodes)
            if d == 0:
                build = torch.zeros(1, dtype=torch.int32, device=self.dev)
                hist = self._build_hist(ridx, seg, build, 1).clone()
            else:
                gc = self._global_counts(counts).contiguous()
                build = torch.empty(nodes // 2, dtype=torch.int32, device=self.dev)
                L.check(lib.kdl_gbdt_choose_build(gc.data_ptr(), int(gc.dtype == torch.float32), nodes,
                                                  build.data_ptr(), L.stream_ptr()), "kdl_gbdt_choose_build")
                built = self._build_hist(ridx, seg, build, nodes // 2)
                # subtraction trick: sibling = parent - built, both children written in one pass
                hist = torch.empty(nodes, self.F, NBINS, 2, dtype=torch.float32, device=self.dev)
                L.check(lib.kdl_gbdt_assemble_hist(parent_hist.data_ptr(), built.data_ptr(), build.data_ptr(), nodes,
                                                   self.F, hist.data_ptr(), L.stream_ptr()), "kdl_gbdt_assemble_hist")
                L.count(2)
            splits = torch.empty(nodes, 8, dtype=torch.float32, device=self.dev)
            L.check(lib.kdl_gbdt_split_eval(hist.data_ptr(), nodes, self.F, self.lam, self.gamma, self.mcw,
                                            splits.data_ptr(), L.stream_ptr()), "kdl_gbdt_split_eval")
            L.check(lib.kdl_gbdt_apply_split(self.node.data_ptr(), self.bins.data_ptr(), FSTRIDE, splits.data_ptr(),
                                             self.n, L.stream_ptr()), "kdl_gbdt_apply_split")
            L.count(2)
            tree.splits.append(splits)
            parent_hist = hist
        # leaf values from the last level's split statistics
        last = tree.splits[-1]
        leaf = torch.empty(1 << self.depth, dtype=torch.float32, device=self.dev)
        L.check(lib.kdl_gbdt_leaf_values(last.data_ptr(), last.shape[0], self.lam, self.lr, leaf.data_ptr(),
                                         L.stream_ptr()), "kdl_gbdt_leaf_values")
        L.count(1)
        tree.leaf_value = leaf
        L.check(lib.kdl_gbdt_update_pred(self.pred.data_ptr(), self.node.data_ptr(), leaf.data_ptr(), self.n,
                                         L.stream_ptr()), "kdl_gbdt_update_pred")
        L.count(1)
        self.trees.append(tree)
        return tree

    def logloss(self) -> torch.Tensor:
        return torch.nn.functional.binary_cross_entropy_with_logits(self.pred, self.y)

    def predict_bins(self, bins: torch.Tensor) -> torch.Tensor:
        """Margin prediction for already-quantised rows (used by tests / evaluation)."""
        n = bins.shape[0]
        out = torch.zeros(n, dtype=torch.float32, device=bins.device)
        for t in self.trees:
            node = torch.zeros(n, dtype=torch.int64, device=bins.device)
            for s in t.splits:
                f = s[node, 0].long()
                thr = s[node, 1].long()
                b = bins[torch.arange(n, device=bins.device), f].long()
                node = 2 * node + (b > thr).long()
            out += t.leaf_value[node]
        return out


# ---------------------------------------------------------------------- model files
def export_model(cuts: torch.Tensor, trees: List[Tree], **meta) -> dict:
    """What a Predict job needs: the cut points that define the bins and, per tree, (feature,
    threshold bin) of every internal node level by level plus the leaf values."""
    return {"format": "kubedl-b200-gbdt-v1", "cuts": cuts.detach().float().cpu(),
            "trees": [{"feat": [s[:, 0].detach().long().cpu() for s in t.splits],
                       "thr": [s[:, 1].detach().long().cpu() for s in t.splits],
                       "leaf": t.leaf_value.detach().float().cpu()} for t in trees],
            "meta": dict(meta)}


def save_model(path: str, model: dict) -> None:
    import os as _os
    _os.makedirs(_os.path.dirname(_os.path.abspath(path)), exist_ok=True)
    tmp = path + ".tmp"
    torch.save(model, tmp)
    _os.replace(tmp, path)


def load_model(path: str) -> dict:
    m = torch.load(path, map_location="cpu", weights_only=False)
    if m.get("format") != "kubedl-b200-gbdt-v1":
        raise ValueError(f"{path} is not a kubedl-b200 GBDT model")
    return m


def predict_margin(model: dict, X: torch.Tensor) -> torch.Tensor:
    """Sum of leaf values over the trees for raw (un-binned) rows, on X's device."""
    dev = X.device
    bins = quantize(X, model["cuts"].to(dev))
    n = X.shape[0]
    rows = torch.arange(n, device=dev)
    out = torch.zeros(n, dtype=torch.float32, device=dev)
    for t in model["trees"]:
        node = torch.zeros(n, dtype=torch.int64, device=dev)
        for feat, thr in zip(t["feat"], t["thr"]):
            f, th = feat.to(dev)[node], thr.to(dev)[node]
            node = 2 * node + (bins[rows, f].long() > th).long()
        out += t["leaf"].to(dev)[node]
    return out


def reference_tree_torch(bins: torch.Tensor, gh: torch.Tensor, F: int, depth: int, lam: float, lr: float,
                         gamma: float = 0.0, mcw: float = 1.0, allreduce=None):
    """Plain PyTorch fp64 oracle of one tree (histogram via index_add, exhaustive gain scan).
    `allreduce(tensor)` (in place, SUM over ranks) makes it the distributed algorithm: every rank
    histograms its own rows, the histograms and leaf sums are summed, and all ranks pick the
    same splits — what XGBoost's `hist` does over Rabit, and what the CUDA trainer fuses."""
    n = bins.shape[0]
    node = torch.zeros(n, dtype=torch.int64, device=bins.device)
    g, h = gh[:, 0].double(), gh[:, 1].double()
    splits = []
    for d in range(depth):
        nodes = 1 << d
        hist = torch.zeros(nodes * F * NBINS, 2, dtype=torch.float64, device=bins.device)
        for f in range(F):
            idx = (node * F + f) * NBINS + bins[:, f].long()
            hist.index_add_(0, idx, torch.stack([g, h], 1))
        if allreduce is not None:
            allreduce(hist)
        hist = hist.view(nodes, F, NBINS, 2)
        cs = hist.cumsum(2)
        tot = cs[:, :, -1:, :]
        GL, HL = cs[..., 0], cs[..., 1]
        GR, HR = tot[..., 0] - GL, tot[..., 1] - HL
        gain = 0.5 * (GL ** 2 / (HL + lam) + GR ** 2 / (HR + lam) - tot[..., 0] ** 2 / (tot[..., 1] + lam)) - gamma
        ok = (HL >= mcw) & (HR >= mcw)
        ok[:, :, -1] = False
        gain = torch.where(ok, gain, torch.full_like(gain, -1.0))
        flat = gain.view(nodes, -1)
        best, code = flat.max(1)
        valid = best > 0
        feat = torch.where(valid, code // NBINS, torch.zeros_like(code))
        thr = torch.where(valid, code % NBINS, torch.full_like(code, 255))
        splits.append((feat, thr, best, valid))
        b = bins[torch.arange(n, device=bins.device), feat[node]].long()
        node = 2 * node + (b > thr[node]).long()
    leaves = 1 << depth
    G = torch.zeros(leaves, dtype=torch.float64, device=bins.device).index_add_(0, node, g)
    H = torch.zeros(leaves, dtype=torch.float64, device=bins.device).index_add_(0, node, h)
    if allreduce is not None:
        allreduce(G)
        allreduce(H)
    return splits, (-G / (H + lam) * lr), node


def tree_from_reference(splits, leaf: torch.Tensor) -> Tree:
    """The oracle's output in the trainer's `Tree` layout (columns 0 / 1 = feature / threshold)."""
    t = Tree(depth=len(splits))
    for feat, thr, gain, valid in splits:
        row = torch.zeros(feat.shape[0], 8, dtype=torch.float32)
        row[:, 0], row[:, 1], row[:, 2], row[:, 7] = feat.float(), thr.float(), gain.float(), valid.float()
        t.splits.append(row)
    t.leaf_value = leaf.float()
    return t
