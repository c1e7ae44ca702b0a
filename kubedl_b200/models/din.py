"""DIN-style sparse CTR model over a row-sharded embedding table (BASELINE.json config #5:
"XDLJob-shaped sparse CTR model (10M-row embedding, DIN-style) with embedding all-to-all
dispatch/combine over NVSwitch at 8 GPUs").

Per sample: user id, candidate item id and L history item ids index ONE table E[V, 64] (fp32
rows, each shard in the symmetric heap).  Rows are INTERLEAVED over the ranks (row id on rank
id % world): NVSwitch gives every peer the same bandwidth, so the only thing placement can get wrong
is load — with row ranges a frequency-sorted (Zipf) vocabulary sends half of all lookups to rank 0
(`sharding="range"` / KUBEDL_DIN_SHARDING=range keeps that layout for A/B runs).  Forward =
`kdl_din_gather_attend` (remote pull + attention pooling + concat, kernel N4) -> MLP 256-512-256-1
on the tcgen05 GEMM -> sigmoid BCE.  Backward = MLP backward on the GEMM (nn / tn variants) ->
device barrier -> `kdl_din_scatter_update` (pooling backward + remote red.add sparse-SGD, kernel N5)
-> fused all-reduce of the dense MLP gradients -> fused SGD.

Ordering of table reads and writes across ranks (synchronous SGD, same result as one GPU holding the
whole table): the flag barrier in front of the scatter makes sure EVERY rank's forward gather of step
N is complete before ANY rank applies step-N updates to the shards (a slow rank would otherwise read
rows its peers are already adding to, 16 bytes at a time); the entry barrier of the dense-gradient
all-reduce orders every rank's step-N updates before anybody's step-N+1 gather.  Both barriers keep
their epoch on the device, so the step stays one replayable CUDA graph.

The XDLJob controller supplies TASK_NAME/TASK_INDEX (reference
controllers/xdl/xdljob_controller.go:191-217); XDL's own PS/worker exchange over ZooKeeper is what
this replaces.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from ..ops import lib as L
from ..ops import nn as K
from ..ops.gemm import Gemm
from .arena import ParamArena

L.register_signatures({
    "kdl_din_gather_attend": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "kdl_din_scatter_update": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_float, C.c_void_p]),
    "kdl_din_bce_grad": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "kdl_relu_bwd_colsum_workspace_bytes": (C.c_size_t, [C.c_int]),
    "kdl_relu_bwd_colsum": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t,
                                      C.c_void_p]),
})

D = 64


class DINModel:
    def __init__(self, vocab: int = 10_000_000, hist_len: int = 62, hidden=(512, 256), lr: float = 0.05,
                 emb_lr: Optional[float] = None, heap=None, group=None, seed: int = 0, device=None,
                 sharding: Optional[str] = None):
        import os as _os
        import torch.distributed as dist
        sharding = sharding or _os.environ.get("KUBEDL_DIN_SHARDING", "interleaved")
        assert sharding in ("interleaved", "range"), sharding
        self.sharding = sharding
        self.interleaved = 1 if sharding == "interleaved" else 0
        self.lib = L.lib()
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.dev = device or torch.device("cuda", torch.cuda.current_device())
        self.vocab, self.L, self.lr = vocab, hist_len, lr
        self.emb_lr = lr if emb_lr is None else emb_lr
        self.rows_per_rank = (vocab + self.world - 1) // self.world
        shard_bytes = self.rows_per_rank * D * 4
        if heap is None:
            from ..parallel.symm import SymmetricHeap
            heap = SymmetricHeap(shard_bytes + (64 << 20), group=group)
        self.heap = heap
        self.shard_off, self.shard = heap.alloc_tensor((self.rows_per_rank, D), torch.float32)
        g = torch.Generator(device=self.dev).manual_seed(seed * 1000 + self.rank)
        self.shard.copy_(torch.randn(self.rows_per_rank, D, device=self.dev, generator=g) * 0.05)
        self._shard_tab = heap.ptr_table(self.shard_off)
        # dense part
        self.arena = ParamArena()
        h1, h2 = hidden
        self.dims = (4 * D, h1, h2, 8)     # the scalar output is padded to 8 columns for the GEMM
        for i, (k_in, k_out) in enumerate(zip(self.dims[:-1], self.dims[1:]), start=1):
            self.arena.declare(f"w{i}", (k_out, k_in), decay=True, init="kaiming", fan=k_in)
            self.arena.declare(f"b{i}", (k_out,), decay=False, init="zeros")
        order = [n for i in (3, 2, 1) for n in (f"w{i}", f"b{i}")]
        self.arena.finalize(self.dev, order, heap=heap, seed=seed)   # same seed -> same MLP on every rank
        with torch.no_grad():
            self.arena.w32("w3")[1:].zero_()
            self.arena.p16.copy_(self.arena.p32)
        self.gemm = Gemm(self.dev)
        # scratch of the fused "ReLU backward + bias gradient" kernel (its counters must start at zero, once)
        self._cs_ws = torch.zeros(self.lib.kdl_relu_bwd_colsum_workspace_bytes(max(self.dims[1:])), dtype=torch.uint8,
                                  device=self.dev)
        self.ar = self.gather_done = None
        if self.world > 1:
            from ..parallel.collectives import FusedAllReduce, SymmBarrier
            self.ar = FusedAllReduce(heap)
            self.gather_done = SymmBarrier(heap)      # its own flag channel
            dist.barrier(group=group)
        self.saved = None

    # ------------------------------------------------------------------ forward / backward
    def forward_z(self, user: torch.Tensor, cand: torch.Tensor, hist: torch.Tensor) -> torch.Tensor:
        """Gather + attention + MLP; returns the padded [B, 8] bf16 output (column 0 is the logit)."""
        B = user.numel()
        x = torch.empty(B, 4 * D, dtype=torch.bfloat16, device=self.dev)
        hs = torch.empty(B, self.L, D, dtype=torch.bfloat16, device=self.dev)
        aw = torch.empty(B, self.L, dtype=torch.float32, device=self.dev)
        L.check(self.lib.kdl_din_gather_attend(self._shard_tab, self.world, self.rows_per_rank, self.interleaved, user.data_ptr(),
                                               cand.data_ptr(), hist.data_ptr(), B, self.L, D, x.data_ptr(),
                                               hs.data_ptr(), aw.data_ptr(), L.stream_ptr()), "kdl_din_gather_attend")
        L.count(1)
        a = self.arena
        h1 = self.gemm.nt(x, a.w16("w1"), bias=a.w16("b1"), relu=True)
        h2 = self.gemm.nt(h1, a.w16("w2"), bias=a.w16("b2"), relu=True)
        z = self.gemm.nt(h2, a.w16("w3"), bias=a.w16("b3"))
        self.saved = (user, cand, hist, x, hs, aw, h1, h2)
        return z

    def forward(self, user: torch.Tensor, cand: torch.Tensor, hist: torch.Tensor) -> torch.Tensor:
        return self.forward_z(user, cand, hist)[:, 0].float()

    def loss_and_dz(self, z: torch.Tensor, label: torch.Tensor):
        """Mean BCE-with-logits of z[:, 0] and its gradient as the padded [B, 8] bf16 matrix — one kernel."""
        B = z.shape[0]
        loss = torch.empty(1, dtype=torch.float32, device=self.dev)
        dz = torch.empty(B, 8, dtype=torch.bfloat16, device=self.dev)
        L.check(self.lib.kdl_din_bce_grad(z.data_ptr(), z.stride(0), label.data_ptr(), B, loss.data_ptr(), dz.data_ptr(),
                                          L.stream_ptr()), "kdl_din_bce_grad")
        L.count(1)
        return loss[0], dz

    def _relu_bwd_bias(self, dh: torch.Tensor, h: Optional[torch.Tensor], dbias: torch.Tensor) -> torch.Tensor:
        """dh *= (h > 0) in place (h None: no mask) and dbias = column sums of dh — one kernel."""
        M, N = dh.shape
        L.check(self.lib.kdl_relu_bwd_colsum(dh.data_ptr(), L.ptr(h), M, N, dbias.data_ptr(), self._cs_ws.data_ptr(),
                                             self._cs_ws.numel(), L.stream_ptr()), "kdl_relu_bwd_colsum")
        L.count(1)
        return dh

    def backward(self, dlogit: torch.Tensor) -> None:
        """dlogit: fp32 [B] gradient of the loss w.r.t. the logits (public API; the training step hands the
        padded bf16 matrix straight from the loss kernel to backward_dz)."""
        dz = torch.zeros(dlogit.numel(), 8, dtype=torch.bfloat16, device=self.dev)
        dz[:, 0] = dlogit.to(torch.bfloat16)
        self.backward_dz(dz)

    def backward_dz(self, dz: torch.Tensor) -> None:
        user, cand, hist, x, hs, aw, h1, h2 = self.saved
        self.saved = None
        a, B = self.arena, user.numel()
        # columns 1..7 of dz are zero, so rows 1..7 of grad(w3) and grad(b3) come out exactly zero by themselves
        self.gemm.tn(dz, h2, out=a.grad("w3"))
        self._relu_bwd_bias(dz, None, a.grad("b3"))
        dh2 = self._relu_bwd_bias(self.gemm.nn(dz, a.w16("w3")), h2, a.grad("b2"))
        self.gemm.tn(dh2, h1, out=a.grad("w2"))
        dh1 = self._relu_bwd_bias(self.gemm.nn(dh2, a.w16("w2")), h1, a.grad("b1"))
        self.gemm.tn(dh1, x, out=a.grad("w1"))
        dx = self.gemm.nn(dh1, a.w16("w1"))
        if self.gather_done is not None:
            # nobody updates a shard while a slower peer may still be gathering this step's rows from it
            self.gather_done()
        L.check(self.lib.kdl_din_scatter_update(self._shard_tab, self.world, self.rows_per_rank, self.interleaved, user.data_ptr(),
                                                cand.data_ptr(), hist.data_ptr(), B, self.L, D, x.data_ptr(),
                                                dx.data_ptr(), hs.data_ptr(), aw.data_ptr(), self.emb_lr,
                                                L.stream_ptr()), "kdl_din_scatter_update")
        L.count(1)

    def step(self, user, cand, hist, label) -> torch.Tensor:
        """One training step.  With a batch of a few thousand samples every kernel is tiny and the step is
        launch-bound (≈ 40 launches), so after one eager step the sequence is captured into a CUDA graph
        that reads the batch from static buffers; KUBEDL_DIN_GRAPH=0 keeps it eager."""
        import os as _os
        if _os.environ.get("KUBEDL_DIN_GRAPH", "1") == "0":
            return self._step_eager(user, cand, hist, label)
        key = (user.numel(), tuple(hist.shape), self.lr, self.emb_lr)    # kernel arguments baked into the graph
        g = getattr(self, "_graphs", None)
        if g is None:
            g = self._graphs = {}
        ent = g.get(key)
        if ent is None:
            g[key] = "warm"
            return self._step_eager(user, cand, hist, label)      # loads kernels, sizes workspaces
        if ent == "warm":
            bufs = (user.clone(), cand.clone(), hist.clone(), label.clone())
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            n0 = L.launch_count()
            with torch.cuda.graph(graph):
                loss = self._step_eager(*bufs)
            ent = g[key] = (graph, bufs, loss, L.launch_count() - n0)
        graph, bufs, loss, launches = ent
        for b, t in zip(bufs, (user, cand, hist, label)):
            b.copy_(t)
        graph.replay()
        L.count(launches)
        return loss.clone()

    def _step_eager(self, user, cand, hist, label) -> torch.Tensor:
        z = self.forward_z(user, cand, hist)
        loss, dz = self.loss_and_dz(z, label)
        self.backward_dz(dz)
        a = self.arena
        if self.ar is not None:
            # dense-gradient all-reduce; its entry barrier also orders every rank's table updates
            # of this step before anybody's gathers of the next step (the gather -> update order inside
            # a step is the SymmBarrier in backward())
            self.ar(a.g16_off, a.total, torch.bfloat16)
        K.sgd_momentum(a.p32, a.mom, a.g16, a.p16, self.lr, 0.0, 0.0, wd_count=a.wd_count)
        return loss

    # ------------------------------------------------------------------ placement
    def owner_of(self, ids: torch.Tensor) -> torch.Tensor:
        """Rank that holds each row id."""
        if self.interleaved:
            return ids % self.world
        return torch.div(ids, self.rows_per_rank, rounding_mode="floor")

    def local_of(self, ids: torch.Tensor) -> torch.Tensor:
        """Row index inside the owner's shard."""
        if self.interleaved:
            return torch.div(ids, self.world, rounding_mode="floor")
        return ids - self.owner_of(ids) * self.rows_per_rank

    def full_table(self, shards) -> torch.Tensor:
        """The [world * rows_per_rank, D] table in id order from the list of per-rank shards."""
        if self.interleaved:
            return torch.stack(list(shards), dim=1).reshape(-1, shards[0].shape[-1])
        return torch.cat(list(shards), 0)

    # ------------------------------------------------------------------ checkpoint (per-rank shard)
    def state_dict(self) -> dict:
        return {"shard": self.shard.detach().cpu(), "arena": self.arena.state_dict(), "rank": self.rank,
                "rows_per_rank": self.rows_per_rank, "sharding": self.sharding, "world": self.world}

    def load_state_dict(self, sd: dict) -> None:
        assert sd["rank"] == self.rank and sd["rows_per_rank"] == self.rows_per_rank
        assert sd.get("sharding", "range") == self.sharding and sd.get("world", self.world) == self.world, \
            "the checkpoint was written with another row placement"
        self.shard.copy_(sd["shard"])
        self.arena.load_state_dict(sd["arena"])


def synthetic_batch(B: int, L: int, vocab: int, device, seed: int = 0, zipf: bool = True):
    """Skewed ids (hot items), variable-length histories padded with -1, a learnable label."""
    g = torch.Generator(device=device).manual_seed(seed)
    def ids(shape):
        if zipf:
            u = torch.rand(shape, device=device, generator=g)
            return (u.pow(3) * vocab).long().clamp_(0, vocab - 1)
        return torch.randint(0, vocab, shape, device=device, generator=g)
    user, cand, hist = ids((B,)), ids((B,)), ids((B, L))
    lens = torch.randint(1, L + 1, (B, 1), device=device, generator=g)
    hist = torch.where(torch.arange(L, device=device)[None, :] < lens, hist, torch.full_like(hist, -1))
    label = (cand % 2).float()          # learnable through the candidate's embedding row
    return user, cand, hist, label
