"""Sharded embedding table over the symmetric heap: the generic dispatch / combine of SURVEY.md §5.8 (`a2a_gather`,
`a2a_scatter_add`) as two kernels that address the owners' shards directly over NVLink (csrc/embedding.cu,
`kdl_emb_gather` / `kdl_emb_scatter_add`).  The DIN model has its own fused variants (gather + attention pooling,
pooling backward + update); this is the building block without a model around it.

    emb = ShardedEmbedding(num_embeddings=10_000_000, dim=64)          # one process per GPU, after init_process_group
    vec = emb.lookup(ids)                  # [n, dim] bf16 (or fp32), rows pulled from whichever rank owns them
    ...                                    # forward / backward of whatever consumes them
    emb.barrier()                          # every rank has finished reading this step's rows
    emb.apply_grad(ids, dvec, lr=0.05)     # E[ids] -= lr * dvec on the owners (duplicates add up)
    emb.barrier()                          # every rank's updates have landed before the next lookup

Rows are interleaved over the ranks by default (row id on rank id % world): NVSwitch gives every peer the same
bandwidth, so only load balance matters, and frequency-sorted vocabularies put their hot rows at low ids.

Stock equivalent: all_to_all_single(ids) + index_select + all_to_all_single(vectors), and all_to_all_single(grads) +
index_add_ (what XDL's parameter servers do for an XDLJob; the reference operator only provides TASK_NAME / TASK_INDEX
/ ZK_ADDR for them, reference controllers/xdl/xdljob_controller.go:191-217).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from ..ops import lib as L

L.register_signatures({
    "kdl_emb_gather": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_int64,
                                 C.c_void_p, C.c_int, C.c_void_p]),
    "kdl_emb_scatter_add": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_int64,
                                      C.c_void_p, C.c_int, C.c_float, C.c_void_p]),
})


class ShardedEmbedding:
    def __init__(self, num_embeddings: int, dim: int, heap=None, group=None, sharding: Optional[str] = None,
                 init_std: float = 0.05, seed: int = 0, device=None):
        import torch.distributed as dist
        assert dim % 4 == 0, "row width must be a multiple of 4 floats"
        self.lib = L.lib()
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.group = group
        self.dev = device or torch.device("cuda", torch.cuda.current_device())
        self.num_embeddings, self.dim = num_embeddings, dim
        sharding = sharding or os.environ.get("KUBEDL_EMB_SHARDING", "interleaved")
        assert sharding in ("interleaved", "range"), sharding
        self.sharding, self.interleaved = sharding, int(sharding == "interleaved")
        self.rows_per_rank = (num_embeddings + self.world - 1) // self.world
        if heap is None:
            from .symm import SymmetricHeap
            heap = SymmetricHeap(self.rows_per_rank * dim * 4 + (8 << 20), group=group)
        self.heap = heap
        self.shard_off, self.shard = heap.alloc_tensor((self.rows_per_rank, dim), torch.float32)
        g = torch.Generator(device=self.dev).manual_seed(seed * 1000 + self.rank)
        self.shard.copy_(torch.randn(self.rows_per_rank, dim, device=self.dev, generator=g) * init_std)
        self._tab = heap.ptr_table(self.shard_off)
        self._barrier = None
        if self.world > 1:
            from .collectives import SymmBarrier
            self._barrier = SymmBarrier(heap, channel=6)
            dist.barrier(group=group)

    # ------------------------------------------------------------------ placement
    def owner_of(self, ids: torch.Tensor) -> torch.Tensor:
        return ids % self.world if self.interleaved else torch.div(ids, self.rows_per_rank, rounding_mode="floor")

    def local_of(self, ids: torch.Tensor) -> torch.Tensor:
        if self.interleaved:
            return torch.div(ids, self.world, rounding_mode="floor")
        return ids - self.owner_of(ids) * self.rows_per_rank

    # ------------------------------------------------------------------ dispatch / combine
    def lookup(self, ids: torch.Tensor, dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
        """out[..., :] = E[ids[...], :]; ids < 0 are padding (zero rows)."""
        assert ids.dtype == torch.int64 and ids.is_cuda and dtype in (torch.bfloat16, torch.float32)
        flat = ids.reshape(-1).contiguous()
        out = torch.empty(flat.numel(), self.dim, dtype=dtype, device=self.dev)
        L.check(self.lib.kdl_emb_gather(self._tab, self.world, self.rows_per_rank, self.interleaved, self.dim,
                                        flat.data_ptr(), flat.numel(), out.data_ptr(), int(dtype == torch.bfloat16),
                                        L.stream_ptr()), "kdl_emb_gather")
        L.count(1)
        return out.view(*ids.shape, self.dim)

    def scatter_add(self, ids: torch.Tensor, grad: torch.Tensor, scale: float = 1.0) -> None:
        """E[ids[i], :] += scale * grad[i, :] on the owning ranks."""
        flat = ids.reshape(-1).contiguous()
        g = grad.reshape(-1, self.dim)
        assert g.shape[0] == flat.numel() and g.is_contiguous() and g.dtype in (torch.bfloat16, torch.float32)
        L.check(self.lib.kdl_emb_scatter_add(self._tab, self.world, self.rows_per_rank, self.interleaved, self.dim,
                                             flat.data_ptr(), flat.numel(), g.data_ptr(), int(g.dtype == torch.bfloat16),
                                             float(scale), L.stream_ptr()), "kdl_emb_scatter_add")
        L.count(1)

    def apply_grad(self, ids: torch.Tensor, grad: torch.Tensor, lr: float) -> None:
        """Sparse SGD: E[ids] -= lr * grad."""
        self.scatter_add(ids, grad, -lr)

    def barrier(self) -> None:
        """Device-side barrier across the ranks (no host round trip; CUDA-graph replayable)."""
        if self._barrier is not None:
            self._barrier()

    # ------------------------------------------------------------------ checkpoint (per-rank shard)
    def state_dict(self) -> dict:
        return {"shard": self.shard.detach().cpu(), "rank": self.rank, "world": self.world, "sharding": self.sharding,
                "rows_per_rank": self.rows_per_rank, "dim": self.dim}

    def load_state_dict(self, sd: dict) -> None:
        assert (sd["rank"], sd["world"], sd["sharding"], sd["rows_per_rank"], sd["dim"]) == \
            (self.rank, self.world, self.sharding, self.rows_per_rank, self.dim), "checkpoint written with another layout"
        self.shard.copy_(sd["shard"])
