"""Fused collectives over the symmetric heap (kernel N1, csrc/allreduce.cu).

`FusedAllReduce` is what the DDP reducer calls per gradient bucket.  It never touches NCCL:
the kernel does barrier -> reduce-scatter -> scale/cast -> all-gather -> barrier over peer
pointers (or NVLS multicast when the heap has a multicast mapping).  `NcclAllReduce` is the
stock path kept for A/B timing only (the baseline worker of BASELINE.md §3).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from ..ops import lib as L
from .symm import FLAG_CHANNELS, SymmetricHeap

_DT = {torch.bfloat16: 0, torch.float32: 1}


class FusedAllReduce:
    def __init__(self, heap: SymmetricHeap, timeout_s: float = 30.0, max_blocks: int = 0,
                 use_nvls: Optional[bool] = None):
        self.heap = heap
        self.lib = L.lib()
        self.timeout_ns = int(timeout_s * 1e9)
        if max_blocks == 0:
            max_blocks = int(os.environ.get("KUBEDL_AR_BLOCKS", "0"))
        self.max_blocks = max_blocks
        if use_nvls is None:
            # measured on 8xB200 (profiles/allreduce_*gpu.log): in-switch reduction wins from
            # 4 ranks up (64 MB: 0.177 ms NVLS vs 0.228 ms P2P at W=8); at W=2 plain P2P is faster
            env = os.environ.get("KUBEDL_NVLS", "auto")
            use_nvls = (heap.world >= 4) if env == "auto" else (env != "0")
        self.use_nvls = bool(use_nvls and heap.mc_ptr)
        self._flag_tab = heap.ptr_table(heap.flags_off)
        self._status_ptr = heap.peer_ptrs[heap.rank] + heap.status_off
        self._tabs: dict = {}

    def _tab(self, off: int):
        t = self._tabs.get(off)
        if t is None:
            t = self.heap.ptr_table(off)
            self._tabs[off] = t
        return t

    def __call__(self, in_off: int, n: int, in_dtype: torch.dtype = torch.bfloat16,
                 out_off: Optional[int] = None, out_dtype: Optional[torch.dtype] = None,
                 scale: Optional[float] = None, channel: int = 0) -> None:
        """All-reduce n elements at heap byte offset `in_off` (in place unless out_off given);
        result = sum * scale (default 1/world)."""
        h = self.heap
        out_off = in_off if out_off is None else out_off
        out_dtype = in_dtype if out_dtype is None else out_dtype
        scale = (1.0 / h.world) if scale is None else scale
        nvls = self.use_nvls and in_dtype == out_dtype
        rc = self.lib.kdl_allreduce_fused(
            self._tab(in_off), self._tab(out_off), self._flag_tab,
            h.mc(in_off) if nvls else None, h.mc(out_off) if nvls else None,
            h.world, h.rank, n, _DT[in_dtype], _DT[out_dtype], scale,
            h.epoch_state.data_ptr(), FLAG_CHANNELS, channel, self.max_blocks, 0,
            self._status_ptr, self.timeout_ns, L.stream_ptr())
        L.check(rc, "kdl_allreduce_fused")
        if h.world > 1 or in_off != out_off or scale != 1.0:
            L.count(1)


class SymmBarrier:
    """Device-side barrier over the symmetric heap's flag words (`kdl_symm_barrier`): what this rank
    enqueued before it has finished on every rank before anything enqueued after it starts anywhere.
    One tiny launch, no host round trip, replayable from a CUDA graph (the epoch lives on the device).
    Every user takes its own flag `channel` (the gradient all-reduce uses the low ones)."""

    def __init__(self, heap: SymmetricHeap, channel: int = FLAG_CHANNELS - 1, timeout_s: float = 30.0):
        assert 0 <= channel < FLAG_CHANNELS
        self.heap, self.channel = heap, channel
        self.lib = L.lib()
        self.timeout_ns = int(timeout_s * 1e9)
        self._flag_tab = heap.ptr_table(heap.flags_off)
        self._status_ptr = heap.peer_ptrs[heap.rank] + heap.status_off

    def __call__(self) -> None:
        h = self.heap
        if h.world == 1:
            return
        rc = self.lib.kdl_symm_barrier(self._flag_tab, h.world, h.rank, h.epoch_state.data_ptr(), FLAG_CHANNELS,
                                       self.channel, 0, self._status_ptr, self.timeout_ns, L.stream_ptr())
        L.check(rc, "kdl_symm_barrier")
        L.count(1)


def emulated_barrier(world: int, device, channel: int = FLAG_CHANNELS - 1):
    """Single-GPU protocol test of `kdl_symm_barrier`: all ranks inside one cooperative launch."""
    lib = L.lib()
    fb = lib.kdl_allreduce_flags_bytes(FLAG_CHANNELS)
    flags = [torch.zeros(fb // 4, dtype=torch.int32, device=device) for _ in range(world)]
    epochs = torch.zeros(world * fb // 4, dtype=torch.int32, device=device)
    status = torch.zeros(64, dtype=torch.int32, device=device)
    ft = (C.c_void_p * 8)()
    for r in range(world):
        ft[r] = flags[r].data_ptr()

    def run():
        rc = lib.kdl_symm_barrier(ft, world, 0, epochs.data_ptr(), FLAG_CHANNELS, channel, 1, status.data_ptr(),
                                  int(10e9), L.stream_ptr())
        L.check(rc, "kdl_symm_barrier(emulate)")
        L.count(1)

    run.state = dict(flags=flags, epochs=epochs, status=status)
    return run


class NcclAllReduce:
    """Baseline: torch.distributed all_reduce (NCCL) followed by the separate scale kernel."""

    def __init__(self, group=None):
        self.group = group

    def __call__(self, t: torch.Tensor, scale: Optional[float] = None) -> None:
        import torch.distributed as dist
        if dist.is_initialized() and dist.get_world_size(self.group) > 1:
            dist.all_reduce(t, group=self.group)
            t.mul_(scale if scale is not None else 1.0 / dist.get_world_size(self.group))


def emulated_allreduce(bufs, out_dtype=None, scale=None, max_blocks: int = 0):
    """Single-GPU protocol test: run all `len(bufs)` ranks of the P2P kernel inside one
    cooperative launch.  bufs: same-shape tensors on one device, reduced IN PLACE (or into new
    tensors when out_dtype differs).  Returns the list of per-rank outputs."""
    lib = L.lib()
    W = len(bufs)
    dev = bufs[0].device
    n = bufs[0].numel()
    in_dtype = bufs[0].dtype
    out_dtype = out_dtype or in_dtype
    outs = bufs if out_dtype == in_dtype else [torch.empty(n, dtype=out_dtype, device=dev) for _ in bufs]
    fb = lib.kdl_allreduce_flags_bytes(FLAG_CHANNELS)
    flags = [torch.zeros(fb // 4, dtype=torch.int32, device=dev) for _ in range(W)]
    epochs = torch.zeros(W * fb // 4, dtype=torch.int32, device=dev)
    status = torch.zeros(64, dtype=torch.int32, device=dev)
    it, ot, ft = (C.c_void_p * 8)(), (C.c_void_p * 8)(), (C.c_void_p * 8)()
    for r in range(W):
        it[r], ot[r], ft[r] = bufs[r].data_ptr(), outs[r].data_ptr(), flags[r].data_ptr()
    state = dict(flags=flags, epochs=epochs, status=status)

    def run():
        rc = lib.kdl_allreduce_fused(it, ot, ft, None, None, W, 0, n, _DT[in_dtype], _DT[out_dtype],
                                     (1.0 / W) if scale is None else scale, epochs.data_ptr(),
                                     FLAG_CHANNELS, 0, max_blocks, 1, status.data_ptr(),
                                     int(10e9), L.stream_ptr())
        L.check(rc, "kdl_allreduce_fused(emulate)")
        L.count(1)
        return outs

    run.state = state
    return run
