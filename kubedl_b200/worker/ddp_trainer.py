"""DDP trainer: the public training API of the PyTorchJob worker runtime.

    trainer = DDPTrainer(batch_per_gpu=256)          # reads RANK/WORLD_SIZE/MASTER_* like torchrun
    loss = trainer.train_step(images_u8_pinned, labels_pinned)   # host tensors in, python float out

One process per GPU.  A step is: H2D copy of the batch (side stream, double buffered) ->
u8 normalise -> ResNet-50 forward/backward on the kubedl_b200 kernels -> per-bucket fused
all-reduce over NVLink peer memory, overlapped with backward -> fused momentum-SGD ->
async D2H of the loss.  After warm-up the device part is replayed from one CUDA graph.

The env contract (MASTER_ADDR, MASTER_PORT, WORLD_SIZE, RANK) is exactly what the PyTorchJob
controller injects (reference controllers/pytorch/pytorchjob_controller.go:180-233); the
stock worker it would launch is torch DDP over NCCL (`backend="nccl"` here, for A/B).
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from ..models.resnet import ResNet50
from ..ops import lib as L
from ..parallel.ddp import BucketReducer


def init_distributed(backend: str = "nccl"):
    """Join the rendezvous the operator wired up (env://).  Returns (rank, world, local_rank)."""
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", os.environ.get("KUBEDL_GPU_INDEX", str(rank))))
    if torch.cuda.is_available():
        torch.cuda.set_device(local_rank % max(torch.cuda.device_count(), 1))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "23456")
        kw = {}
        if backend == "nccl" and torch.cuda.is_available():
            kw["device_id"] = torch.device("cuda", torch.cuda.current_device())
        if os.environ.get("KUBEDL_RENDEZVOUS_TIMEOUT"):
            # bound the wait for peers that never come (docs/fault_tolerance.md); unset = torch's default
            import datetime
            kw["timeout"] = datetime.timedelta(seconds=float(os.environ["KUBEDL_RENDEZVOUS_TIMEOUT"]))
        if backend == "gloo" and os.environ["MASTER_ADDR"] in ("127.0.0.1", "localhost", "::1"):
            # all pods of a job share the box (the node agent rewrites service names to loopback): skip gloo's
            # resolution of the container hostname, which costs ~0.5 s and may not resolve at all
            os.environ.setdefault("GLOO_SOCKET_IFNAME", "lo")
        dist.init_process_group(backend=backend, init_method="env://", rank=rank, world_size=world, **kw)
    return rank, world, local_rank


class DDPTrainer:
    def __init__(self, batch_per_gpu: int = 256, lr: float = 0.1, momentum: float = 0.9,
                 weight_decay: float = 1e-4, comm: str = "fused", use_graph: bool = True,
                 use_gemm: Optional[bool] = None, bucket_mb: float = 64.0, image_size: int = 224,
                 num_classes: int = 1000, seed: int = 0):
        import time as _time
        import torch.distributed as dist
        _t = _time.time()
        self.startup = {}          # seconds per start-up phase (time to first step is what a user waits for)

        def _mark(name):
            nonlocal _t
            now = _time.time()
            self.startup[name] = round(now - _t, 3)
            _t = now
        # torch.distributed is only plumbing on the fused path (handle exchange, barriers at start-up and
        # capture): a gloo group joins in a fraction of the time NCCL needs to build a communicator that would
        # never carry a byte (3.6 s of the 6.2 s to the first step at 2 pods).  NCCL is created when it is the
        # data path (--comm nccl) or asked for (KUBEDL_PLUMBING_BACKEND=nccl).
        plumbing = os.environ.get("KUBEDL_PLUMBING_BACKEND", "gloo" if comm == "fused" else "nccl")
        self.rank, self.world, self.local_rank = init_distributed(plumbing)
        _mark("rendezvous")
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.batch, self.lr, self.momentum, self.wd = batch_per_gpu, lr, momentum, weight_decay
        self.image_size = image_size
        torch.backends.cudnn.benchmark = True
        L.lib()  # fail loudly now if the kernel library is missing
        _mark("kernel_lib")
        self.heap = None
        self.comm = comm if self.world > 1 else "none"
        if self.comm == "fused":
            from ..parallel.symm import SymmetricHeap
            # ResNet-50: 25.6 M params -> 51 MB of bf16 gradients (+ flags); 128 MB is ample
            self.heap = SymmetricHeap(128 << 20)
        _mark("symmetric_heap")
        gemm = None
        if use_gemm is None:
            use_gemm = os.environ.get("KUBEDL_GEMM", "1") != "0"
        if use_gemm:
            try:
                from ..ops import gemm as G
                gemm = G.Gemm(self.device)
            except ImportError:
                gemm = None
        self.model = ResNet50(self.device, num_classes=num_classes, heap=self.heap, gemm=gemm, seed=seed)
        # One bucket by default (ResNet-50's 51 MB of bf16 gradients fit in 64 MB): on NVLink the all-reduce takes
        # ~0.15 ms, and overlapping it with backward costs MORE than it hides, because the persistent GEMM /
        # convolution kernels own whole SMs — a bucket's blocks delay the next such kernel by their whole
        # duration, barrier skew included.  Measured (ms/step, 16 MB vs 64 MB buckets): 19.37 / 19.23 at 2 GPUs,
        # 19.57 / 19.45 at 8 (profiles/r2_bucket_size.md).  KUBEDL_BUCKET_MB restores the overlapped schedule.
        bucket_mb = float(os.environ.get("KUBEDL_BUCKET_MB", bucket_mb))
        self.model.reducer = BucketReducer(self.model.arena, heap=self.heap, bucket_mb=bucket_mb,
                                           backend="fused" if self.comm == "fused" else "nccl",
                                           world=self.world)
        if self.world > 1:
            # identical initial weights on every rank (same seed) — verify cheaply
            chk = self.model.arena.p32[:1024].sum().reshape(1).clone()
            if dist.get_backend() == "gloo":
                chk = chk.cpu()
            lst = [torch.empty_like(chk) for _ in range(self.world)]
            dist.all_gather(lst, chk)
            assert all(torch.equal(lst[0], t) for t in lst), "ranks disagree on initial weights"
        _mark("model")
        S = image_size
        self.images = [torch.zeros(batch_per_gpu, S, S, 3, dtype=torch.uint8, device=self.device)
                       for _ in range(2)]
        self.labels = [torch.zeros(batch_per_gpu, dtype=torch.int64, device=self.device) for _ in range(2)]
        self.loss_dev = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
        self.copy_stream = torch.cuda.Stream()
        self._copied = [torch.cuda.Event() for _ in range(2)]
        self._consumed = [torch.cuda.Event() for _ in range(2)]
        self._loss_ready = torch.cuda.Event()
        self.use_graph = use_graph
        self.graphs = [None, None]
        self.launches_per_step = 0
        self.steps_done = 0
        self._slot = 0
        self._pending_loss = False

    # ------------------------------------------------------------------ device step
    def _device_step(self, slot: int) -> None:
        loss = self.model.loss_and_grad(self.images[slot], self.labels[slot])
        self.model.sgd_step(self.lr, self.momentum, self.wd)
        self.loss_dev.copy_(loss)

    def warmup(self, steps: int = 3) -> None:
        """Eager steps (cuDNN autotune, workspace sizing), then graph capture per input slot."""
        for _ in range(max(steps, 1)):
            before = L.launch_count()
            self._device_step(0)
            self.launches_per_step = L.launch_count() - before
        torch.cuda.synchronize()
        if self.use_graph:
            self.capture()

    def capture(self) -> None:
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        for slot in range(2):
            g = torch.cuda.CUDAGraph()
            torch.cuda.synchronize()
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                self._device_step(slot)
            self.graphs[slot] = g
        torch.cuda.synchronize()

    def run_device_step(self, slot: int = 0) -> None:
        """One training step on data already resident in device slot `slot`."""
        if self.graphs[slot] is not None:
            self.graphs[slot].replay()
            L.count(self.launches_per_step)
        else:
            self._device_step(slot)
        self.steps_done += 1

    # ------------------------------------------------------------------ end-to-end step
    def train_step(self, images_u8: torch.Tensor, labels: torch.Tensor, sync: bool = False) -> float:
        """images_u8: pinned host uint8 [B,H,W,3]; labels: pinned host int64 [B].  Returns the loss
        of the PREVIOUS step (pipelined read-back; the current one if sync=True)."""
        slot = self._slot
        cur = torch.cuda.current_stream()
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self._consumed[slot])  # slot free again
            self.images[slot].copy_(images_u8, non_blocking=True)
            self.labels[slot].copy_(labels, non_blocking=True)
            self._copied[slot].record(self.copy_stream)
        prev = float("nan")
        if self._pending_loss:
            self._loss_ready.synchronize()
            prev = float(self.loss_host[0])
        cur.wait_event(self._copied[slot])
        self.run_device_step(slot)
        self._consumed[slot].record(cur)
        self.loss_host.copy_(self.loss_dev, non_blocking=True)
        self._loss_ready.record(cur)
        self._pending_loss = True
        self._slot ^= 1
        if sync:
            self._loss_ready.synchronize()
            return float(self.loss_host[0])
        return prev

    def last_loss(self) -> float:
        self._loss_ready.synchronize()
        return float(self.loss_host[0])

    # ------------------------------------------------------------------ checkpoint / resume
    def state_dict(self) -> dict:
        bn = {}
        mods = [self.model.stem_bn] + [m for b in self.model.blocks for m in b.modules() if hasattr(m, "running_mean")]
        for m in mods:
            bn[m.name] = (m.running_mean.cpu(), m.running_var.cpu())
        return {"arena": self.model.arena.state_dict(), "bn": bn, "step": self.steps_done}

    def load_state_dict(self, sd: dict) -> None:
        self.model.arena.load_state_dict(sd["arena"])
        mods = [self.model.stem_bn] + [m for b in self.model.blocks for m in b.modules() if hasattr(m, "running_mean")]
        for m in mods:
            rm, rv = sd["bn"][m.name]
            m.running_mean.copy_(rm)
            m.running_var.copy_(rv)
        self.steps_done = sd["step"]

    def save(self, path: str) -> None:
        """Rank 0 writes model+optimizer+step; a restarted pod with the same RANK resumes."""
        if self.rank == 0:
            tmp = path + ".tmp"
            torch.save(self.state_dict(), tmp)
            os.replace(tmp, path)

    def load(self, path: str) -> bool:
        if not os.path.exists(path):
            return False
        self.load_state_dict(torch.load(path, map_location="cpu", weights_only=False))
        return True
