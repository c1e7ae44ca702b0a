"""Bucketed gradient reducer (the DDP half of SURVEY.md §2.4 rows N1/N2).

The gradient arena is laid out in backward order, so "gradients ready up to element X" is a
monotonically growing prefix.  Whenever the prefix covers a whole bucket the reducer records
an event on the compute stream and launches the fused all-reduce kernel for that bucket on a
side stream: the collective for bucket k overlaps the backward kernels of bucket k+1.  With
`backend="nccl"` the same schedule drives torch.distributed (baseline for A/B).

Stock equivalent: torch DDP's C++ reducer with autograd hooks + ncclAllReduce per 25 MB bucket,
configured purely by the env the operator injects (reference
controllers/pytorch/pytorchjob_controller.go:180-233).
"""
from __future__ import annotations

from typing import List, Optional

import torch

from .collectives import FusedAllReduce, NcclAllReduce


class BucketReducer:
    def __init__(self, arena, heap=None, bucket_mb: float = 16.0, backend: str = "fused",
                 group=None, world: int = 1):
        self.arena = arena
        self.heap = heap
        self.backend = backend
        self.world = world
        n = arena.total
        per = max(int(bucket_mb * (1 << 20) / 2) // 512 * 512, 512)
        self.bounds: List[int] = []  # bucket k = [bounds[k], bounds[k+1])
        off = 0
        while off < arena.wd_count:
            self.bounds.append(off)
            off = min(off + per, arena.wd_count)
        # the tail of the decayed region and the whole no-decay region form the final bucket
        if not self.bounds:
            self.bounds = [0]
        elif arena.wd_count - self.bounds[-1] < per // 4 and len(self.bounds) > 1:
            self.bounds.pop()
        self.bounds.append(n)
        self.nbuckets = len(self.bounds) - 1
        self.next_bucket = 0
        self.comm_stream = torch.cuda.Stream() if (world > 1) else None
        self.fused = FusedAllReduce(heap) if (backend == "fused" and heap is not None) else None
        self.nccl = NcclAllReduce(group) if backend == "nccl" else None
        self._ev = [torch.cuda.Event() for _ in range(self.nbuckets + 1)]

    def start(self) -> None:
        self.next_bucket = 0

    def _launch(self, k: int) -> None:
        lo, hi = self.bounds[k], self.bounds[k + 1]
        if self.world == 1:
            return
        cur = torch.cuda.current_stream()
        self._ev[k].record(cur)
        with torch.cuda.stream(self.comm_stream):
            self.comm_stream.wait_event(self._ev[k])
            if self.fused is not None:
                self.fused(self.arena.g16_off + lo * 2, hi - lo, torch.bfloat16,
                           channel=k % 8)
            else:
                self.nccl(self.arena.g16[lo:hi])

    def ready_upto(self, end_elem: int) -> None:
        """Backward has produced every gradient in [0, end_elem)."""
        while self.next_bucket < self.nbuckets - 1 and self.bounds[self.next_bucket + 1] <= end_elem:
            self._launch(self.next_bucket)
            self.next_bucket += 1

    def finish(self) -> None:
        """All gradients are ready: flush remaining buckets and join the comm stream."""
        while self.next_bucket < self.nbuckets:
            self._launch(self.next_bucket)
            self.next_bucket += 1
        if self.comm_stream is not None:
            self._ev[self.nbuckets].record(self.comm_stream)
            torch.cuda.current_stream().wait_event(self._ev[self.nbuckets])
