"""Flat parameter / gradient arenas for the DDP worker.

All parameters of a model live in four flat buffers with identical layout:
  p32  fp32 master weights          mom  fp32 momentum
  p16  bf16 compute copy            g16  bf16 gradients (inside the symmetric heap, so the
                                         fused all-reduce works on it in place)
Layout = [decayed weights, in BACKWARD order][non-decayed params (BN scale/shift, biases)], so
gradient buckets complete as a growing prefix while backward runs and one `wd_count` index
separates the two weight-decay regimes for the fused SGD kernel.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import torch


@dataclass
class ParamSpec:
    name: str
    shape: Tuple[int, ...]
    decay: bool
    init: str  # "kaiming" | "ones" | "zeros" | "fc"
    fan: int = 0
    offset: int = 0
    numel: int = 0


class ParamArena:
    ALIGN = 64  # elements; keeps every parameter 128-byte aligned in bf16

    def __init__(self):
        self.specs: List[ParamSpec] = []
        self.by_name: Dict[str, ParamSpec] = {}
        self.total = 0
        self.wd_count = 0
        self.p32 = self.p16 = self.g16 = self.mom = None

    def declare(self, name: str, shape, decay: bool, init: str, fan: int = 0) -> None:
        n = 1
        for s in shape:
            n *= s
        sp = ParamSpec(name, tuple(shape), decay, init, fan, 0, n)
        self.specs.append(sp)
        self.by_name[name] = sp

    def finalize(self, device, backward_order: List[str], heap=None, seed: int = 0) -> None:
        """Assign offsets (decayed params in `backward_order`, then the rest) and allocate."""
        order = [self.by_name[n] for n in backward_order if self.by_name[n].decay]
        rest = [self.by_name[n] for n in backward_order if not self.by_name[n].decay]
        missing = set(self.by_name) - set(backward_order)
        assert not missing, f"params missing from backward order: {missing}"
        off = 0
        for sp in order:
            sp.offset = off
            off += (sp.numel + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        self.wd_count = off
        for sp in rest:
            sp.offset = off
            off += (sp.numel + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        self.total = off
        self.layout = order + rest
        self.p32 = torch.zeros(off, dtype=torch.float32, device=device)
        self.mom = torch.zeros(off, dtype=torch.float32, device=device)
        self.p16 = torch.zeros(off, dtype=torch.bfloat16, device=device)
        if heap is not None:
            self.g16_off, self.g16 = heap.alloc_tensor((off,), torch.bfloat16)
            self.g16.zero_()
        else:
            self.g16_off, self.g16 = None, torch.zeros(off, dtype=torch.bfloat16, device=device)
        g = torch.Generator(device="cpu").manual_seed(seed)
        for sp in self.layout:
            v = self.p32[sp.offset:sp.offset + sp.numel]
            if sp.init == "kaiming":
                std = (2.0 / max(sp.fan, 1)) ** 0.5
                v.copy_(torch.randn(sp.numel, generator=g) * std)
            elif sp.init == "fc":
                bound = 1.0 / max(sp.fan, 1) ** 0.5
                v.copy_((torch.rand(sp.numel, generator=g) * 2 - 1) * bound)
            elif sp.init == "ones":
                v.fill_(1.0)
        self.p16.copy_(self.p32)

    def view(self, buf: torch.Tensor, name: str) -> torch.Tensor:
        sp = self.by_name[name]
        return buf[sp.offset:sp.offset + sp.numel].view(*sp.shape)

    def w16(self, name): return self.view(self.p16, name)
    def w32(self, name): return self.view(self.p32, name)
    def grad(self, name): return self.view(self.g16, name)

    def end_offset(self, name: str) -> int:
        sp = self.by_name[name]
        return sp.offset + (sp.numel + self.ALIGN - 1) // self.ALIGN * self.ALIGN

    def state_dict(self) -> dict:
        return {"p32": self.p32.detach().cpu(), "mom": self.mom.detach().cpu(),
                "layout": [(s.name, s.shape, s.offset) for s in self.layout]}

    def load_state_dict(self, sd: dict) -> None:
        assert [(s.name, tuple(s.shape), s.offset) for s in self.layout] == \
            [(n, tuple(sh), o) for n, sh, o in sd["layout"]], "arena layout mismatch"
        self.p32.copy_(sd["p32"])
        self.mom.copy_(sd["mom"])
        self.p16.copy_(self.p32)
