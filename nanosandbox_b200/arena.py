"""Flat parameter / gradient arenas.

Data layout in HBM: every parameter lives in ONE contiguous fp32 buffer, in
the order in which backward finishes their gradients (ln_f, blocks L-1..0, wpe, tied wte last),
each tensor padded to a multiple of 1024 elements.  The gradient arena, the AdamW moments and the
bf16 shadow copy used as GEMM operands share that layout, so that

  * AdamW is one kernel launch over the whole model (a per-1024-chunk flag selects weight decay),
  * the global grad norm is one reduction over one buffer,
  * a data-parallel gradient bucket is a contiguous slice [start, end) of the gradient arena that
    becomes final at a known point of backward (no gather/scatter copies, unlike DDP's bucket
    views: torch/nn/parallel/distributed.py:667, 803).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Tuple

import torch

CHUNK = 1024


@dataclass
class Slot:
    name: str
    shape: Tuple[int, ...]
    offset: int
    numel: int
    decay: bool


class ParamArena:
    def __init__(self, named: List[Tuple[str, torch.Tensor]], device, with_lowp: bool = True):
        self.slots: Dict[str, Slot] = {}
        off = 0
        for name, t in named:
            n = t.numel()
            self.slots[name] = Slot(name, tuple(t.shape), off, n, t.dim() >= 2)
            off += (n + CHUNK - 1) // CHUNK * CHUNK
        self.numel = off
        self.device = torch.device(device)
        self.params = torch.zeros(off, dtype=torch.float32, device=device)
        self.grads = torch.zeros(off, dtype=torch.float32, device=device)
        self.lowp = torch.zeros(off, dtype=torch.bfloat16, device=device) if with_lowp else None
        flags = torch.zeros(off // CHUNK, dtype=torch.uint8)
        for s in self.slots.values():
            if s.decay:
                flags[s.offset // CHUNK:(s.offset + s.numel + CHUNK - 1) // CHUNK] = 1
        self.chunk_decay = flags.to(device)
        self.lowp_version = -1  # params._version the bf16 shadow was last refreshed at

    def view(self, buf: torch.Tensor, name: str) -> torch.Tensor:
        s = self.slots[name]
        return buf[s.offset:s.offset + s.numel].view(s.shape)

    def p(self, name):
        return self.view(self.params, name)

    def g(self, name):
        return self.view(self.grads, name)

    def w(self, name, lowp: bool):
        """GEMM operand view of a weight: bf16 shadow or the fp32 master."""
        return self.view(self.lowp if lowp else self.params, name)

    def span(self, names: List[str]) -> Tuple[int, int]:
        """[start, end) element range covering `names` (must be adjacent in the arena)."""
        ss = [self.slots[n] for n in names]
        start = min(s.offset for s in ss)
        end = max((s.offset + s.numel + CHUNK - 1) // CHUNK * CHUNK for s in ss)
        return start, end


def ready_order(named: List[Tuple[str, torch.Tensor]], n_layer: int) -> List[Tuple[str, torch.Tensor]]:
    """Re-order nanoGPT's named_parameters() into gradient-ready order of our backward."""
    d = dict(named)
    out: List[Tuple[str, torch.Tensor]] = []

    def take(prefix):
        for suffix in ("weight", "bias"):
            k = f"{prefix}.{suffix}"
            if k in d:
                out.append((k, d.pop(k)))

    take("transformer.ln_f")
    for l in range(n_layer - 1, -1, -1):
        for sub in ("mlp.c_proj", "mlp.c_fc", "ln_2", "attn.c_proj", "attn.c_attn", "ln_1"):
            take(f"transformer.h.{l}.{sub}")
    take("transformer.wpe")
    take("transformer.wte")
    assert not d, f"parameters not placed in the arena: {list(d)}"
    return out


def bucket_ranges(arena: ParamArena, n_layer: int, has_bias: bool) -> List[Tuple[int, int]]:
    """One bucket per transformer block in backward order (ln_f rides with the last block), plus a
    final bucket for wpe + the tied wte, whose gradient is complete only after the embedding
    scatter-add at the very end of backward."""
    sfx = ("weight", "bias") if has_bias else ("weight",)
    ranges = []
    for i, l in enumerate(range(n_layer - 1, -1, -1)):
        names = [f"transformer.h.{l}.{sub}.{s}" for sub in
                 ("mlp.c_proj", "mlp.c_fc", "ln_2", "attn.c_proj", "attn.c_attn", "ln_1") for s in sfx]
        if i == 0:
            names += [f"transformer.ln_f.{s}" for s in sfx]
        ranges.append(arena.span(names))
    ranges.append(arena.span(["transformer.wpe.weight", "transformer.wte.weight"]))
    # the spans tile the arena exactly
    assert ranges[0][0] == 0 and ranges[-1][1] == arena.numel
    for a, b in zip(ranges, ranges[1:]):
        assert a[1] == b[0]
    return ranges
