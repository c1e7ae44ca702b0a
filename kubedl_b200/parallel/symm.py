"""Symmetric heap over NVLink peer memory.

One process per GPU.  Every rank allocates a heap of the same size; handles are exchanged
through the job's rendezvous (the `torch.distributed` group that `MASTER_ADDR/MASTER_PORT/
RANK/WORLD_SIZE` bootstraps — the env contract the operator injects, reference
controllers/pytorch/pytorchjob_controller.go:180-233), after which every rank holds a device
pointer to every peer's heap.  Kernels then load/store peer memory directly.

Two transports:
  * "vmm": cuMemCreate + POSIX fds passed over AF_UNIX sockets; allows binding the heap to
    an NVLS multicast object (multimem.* instructions reduce inside the NVSwitch).
  * "ipc": cudaMalloc + cudaIpcMemHandle (no multicast).
"auto" tries vmm (+multicast) and falls back to ipc.
"""
from __future__ import annotations

import ctypes as C
import os
import socket
import struct
import threading
import time
from typing import List, Optional

import torch

from ..ops import lib as L

FLAG_CHANNELS = 8


class _Holder:
    """Exposes raw device memory to torch through __cuda_array_interface__ (zero copy)."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2, "strides": None,
        }
        self._owner = owner  # keeps the heap alive while tensors exist


def _exchange_fds(rank: int, world: int, fd_by_dst: dict, tag: str, group) -> dict:
    """All-to-all of file descriptors.  fd_by_dst[dst] = fd to send; returns {src: fd}."""
    import torch.distributed as dist
    nonce = [None]
    if rank == 0:
        nonce[0] = f"{os.getpid()}-{time.time_ns()}"
    dist.broadcast_object_list(nonce, src=0, group=group)
    name = lambda r: f"\0kdl-symm-{nonce[0]}-{tag}-{r}"
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(name(rank))
    srv.listen(world)
    expect = sum(1 for src in range(world) if src != rank)
    got: dict = {}
    err: list = []

    def serve():
        try:
            srv.settimeout(60)
            for _ in range(expect):
                conn, _ = srv.accept()
                with conn:
                    msg, fds, _, _ = socket.recv_fds(conn, 16, 1)
                    src = struct.unpack("i", msg[:4])[0]
                    got[src] = fds[0]
        except Exception as e:  # pragma: no cover - GPU box only
            err.append(e)

    th = threading.Thread(target=serve, daemon=True)
    th.start()
    dist.barrier(group=group)  # every listener is up
    for dst, fd in fd_by_dst.items():
        if dst == rank:
            continue
        c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        c.connect(name(dst))
        socket.send_fds(c, [struct.pack("i", rank)], [fd])
        c.close()
    th.join(90)
    srv.close()
    if err:
        raise err[0]
    if len(got) != expect:
        raise RuntimeError(f"fd exchange incomplete: {len(got)}/{expect}")
    return got


class SymmetricHeap:
    def __init__(self, nbytes: int, group=None, mode: str = "auto", device: Optional[int] = None,
                 multicast: bool = True):
        import torch.distributed as dist
        self.lib = L.lib()
        self.group = group
        if dist.is_available() and dist.is_initialized():
            self.rank = dist.get_rank(group)
            self.world = dist.get_world_size(group)
        else:
            self.rank, self.world = 0, 1
        if self.world > 8:
            raise ValueError("SymmetricHeap supports one NVSwitch node (world <= 8)")
        self.device = torch.cuda.current_device() if device is None else device
        self.mode = mode if mode != "auto" else os.environ.get("KUBEDL_SYMM_MODE", "auto")
        self.mc_ptr = 0
        self._mc_handle = 0
        self._handles: List[int] = [0] * self.world
        self.peer_ptrs: List[int] = [0] * self.world
        self.transport = None
        gran = 2 << 20
        if self.world > 1 and self.mode in ("auto", "vmm"):
            g = self.lib.kdl_vmm_granularity(self.device, self.world if multicast else 0)
            if g:
                gran = max(gran, g)
        self.nbytes = (nbytes + gran - 1) // gran * gran
        if self.world == 1:
            self._alloc_local_only()
        else:
            ok = False
            if self.mode in ("auto", "vmm"):
                try:
                    ok = self._init_vmm(multicast)
                except Exception as e:  # pragma: no cover - GPU box only
                    if self.mode == "vmm":
                        raise
                    if self.rank == 0:
                        print(f"[kubedl_b200.symm] vmm transport unavailable ({e}); using cudaIpc", flush=True)
                    ok = False
                # all ranks must agree
                flag = [ok]
                allf = [None] * self.world
                dist.all_gather_object(allf, ok, group=group)
                if not all(allf):
                    if ok:
                        self._release_vmm()
                    ok = False
            if not ok:
                self._init_ipc()
            dist.barrier(group=group)
        self._bump = 0
        self._tensor = torch.as_tensor(_Holder(self.peer_ptrs[self.rank], self.nbytes, self),
                                       device=torch.device("cuda", self.device))
        # reserved head: flags for the collectives + a status word
        self.flags_off = self.alloc(self.lib.kdl_allreduce_flags_bytes(FLAG_CHANNELS))
        self.status_off = self.alloc(256)
        self.epoch_state = torch.zeros(self.lib.kdl_allreduce_flags_bytes(FLAG_CHANNELS) // 4,
                                       dtype=torch.int32, device=torch.device("cuda", self.device))

    # ------------------------------------------------------------------ transports
    def _alloc_local_only(self):
        p = C.c_void_p()
        L.check(self.lib.kdl_symm_malloc(self.nbytes, C.byref(p)), "kdl_symm_malloc")
        self.peer_ptrs[0] = p.value
        self.transport = "local"

    def _init_ipc(self):
        import torch.distributed as dist
        p = C.c_void_p()
        L.check(self.lib.kdl_symm_malloc(self.nbytes, C.byref(p)), "kdl_symm_malloc")
        h = C.create_string_buffer(64)
        L.check(self.lib.kdl_ipc_get_handle(p, h), "kdl_ipc_get_handle")
        handles = [None] * self.world
        dist.all_gather_object(handles, (self.device, h.raw), group=self.group)
        for r, (dev, raw) in enumerate(handles):
            if r == self.rank:
                self.peer_ptrs[r] = p.value
                continue
            q = C.c_void_p()
            L.check(self.lib.kdl_ipc_open_handle(C.create_string_buffer(raw, 64), C.byref(q)),
                    "kdl_ipc_open_handle")
            self.peer_ptrs[r] = q.value
        self.transport = "ipc"

    def _init_vmm(self, multicast: bool) -> bool:
        import torch.distributed as dist
        fd_ok, mc_ok = C.c_int(0), C.c_int(0)
        self.lib.kdl_vmm_supported(self.device, C.byref(fd_ok), C.byref(mc_ok))
        if not fd_ok.value:
            raise RuntimeError("POSIX fd handles unsupported")
        h, p, fd = C.c_ulonglong(), C.c_void_p(), C.c_int(-1)
        L.check(self.lib.kdl_vmm_alloc(self.device, self.nbytes, C.byref(h), C.byref(p), C.byref(fd)),
                "kdl_vmm_alloc")
        self._handles[self.rank] = h.value
        self.peer_ptrs[self.rank] = p.value
        got = _exchange_fds(self.rank, self.world, {d: fd.value for d in range(self.world)}, "mem",
                            self.group)
        os.close(fd.value)
        for src, rfd in got.items():
            hh, pp = C.c_ulonglong(), C.c_void_p()
            L.check(self.lib.kdl_vmm_import(self.device, rfd, self.nbytes, C.byref(hh), C.byref(pp)),
                    "kdl_vmm_import")
            os.close(rfd)
            self._handles[src] = hh.value
            self.peer_ptrs[src] = pp.value
        self.transport = "vmm"
        if multicast and mc_ok.value and os.environ.get("KUBEDL_NVLS", "1") != "0":
            try:
                self._init_multicast()
            except Exception as e:  # pragma: no cover - GPU box only
                if self.rank == 0:
                    print(f"[kubedl_b200.symm] NVLS multicast unavailable ({e}); P2P only", flush=True)
                self.mc_ptr = 0
            oks = [None] * self.world
            dist.all_gather_object(oks, bool(self.mc_ptr), group=self.group)
            if not all(oks):
                self.mc_ptr = 0
        return True

    def _init_multicast(self):
        """Rank 0 creates the multicast object and ships its fd; everyone adds its device,
        binds its own heap allocation and maps the multicast VA."""
        import torch.distributed as dist
        mc = C.c_ulonglong()
        fd = C.c_int(-1)
        if self.rank == 0:
            L.check(self.lib.kdl_mc_create(self.world, self.nbytes, C.byref(mc), C.byref(fd)),
                    "kdl_mc_create")
            send = {d: fd.value for d in range(self.world)}
            null = -1
        else:
            # the exchanger is symmetric (world-1 inbound fds per rank): non-root ranks send a
            # placeholder fd so the accounting stays uniform
            null = os.open(os.devnull, os.O_RDONLY)
            send = {d: null for d in range(self.world)}
        got = _exchange_fds(self.rank, self.world, send, "mc", self.group)
        if self.rank == 0:
            os.close(fd.value)
        else:
            os.close(null)
            L.check(self.lib.kdl_mc_import(got[0], C.byref(mc)), "kdl_mc_import")
        for f in got.values():
            os.close(f)
        self._mc_handle = mc.value
        L.check(self.lib.kdl_mc_add_device(mc.value, self.device), "kdl_mc_add_device")
        dist.barrier(group=self.group)  # every device added before any bind
        q = C.c_void_p()
        L.check(self.lib.kdl_mc_bind_and_map(mc.value, self._handles[self.rank], self.device,
                                             self.nbytes, C.byref(q)), "kdl_mc_bind_and_map")
        self.mc_ptr = q.value
        dist.barrier(group=self.group)

    def _release_vmm(self):
        for r in range(self.world):
            if self._handles[r]:
                self.lib.kdl_vmm_release(self._handles[r], self.peer_ptrs[r], self.nbytes)
                self._handles[r] = 0
                self.peer_ptrs[r] = 0

    # ------------------------------------------------------------------ allocation / views
    def alloc(self, nbytes: int, align: int = 256) -> int:
        off = (self._bump + align - 1) // align * align
        if off + nbytes > self.nbytes:
            raise MemoryError(f"symmetric heap exhausted ({off + nbytes} > {self.nbytes})")
        self._bump = off + nbytes
        return off

    def tensor(self, offset: int, shape, dtype: torch.dtype) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= s
        nb = n * torch.empty((), dtype=dtype).element_size()
        return self._tensor[offset:offset + nb].view(dtype).view(*shape)

    def alloc_tensor(self, shape, dtype: torch.dtype):
        n = 1
        for s in shape:
            n *= s
        off = self.alloc(n * torch.empty((), dtype=dtype).element_size())
        return off, self.tensor(off, shape, dtype)

    def ptr_table(self, offset: int):
        arr = (C.c_void_p * 8)()
        for r in range(self.world):
            arr[r] = self.peer_ptrs[r] + offset
        return arr

    def mc(self, offset: int) -> Optional[int]:
        return (self.mc_ptr + offset) if self.mc_ptr else None

    def status_word(self) -> int:
        return int(self.tensor(self.status_off, (1,), torch.int32).item())
