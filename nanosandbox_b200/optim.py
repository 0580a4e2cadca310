"""Fused AdamW + global-norm clipping over the flat arenas.

Replaces ``torch.optim.AdamW(fused=True)`` as returned by nanoGPT's ``configure_optimizers``
(arithmetic: torch/optim/adam.py:417-419, 457, 476, 531-547) and ``clip_grad_norm_``
(torch/nn/utils/clip_grad.py:96-106, 165-174).  One kernel launch updates every parameter; the
clip coefficient is computed on the device and applied inside that launch (no extra pass over
the gradients, no host sync).
"""
from __future__ import annotations

import math

import torch

from . import ops
from ._lib import NsbError
from .arena import ParamArena


class FusedAdamW(torch.optim.Optimizer):
    def __init__(self, param_groups, arena: ParamArena, lr=1e-3, betas=(0.9, 0.999), eps=1e-8,
                 weight_decay=1e-2):
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(param_groups, defaults)
        self.arena = arena
        dev = arena.device
        self.exp_avg = torch.zeros(arena.numel, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(arena.numel, dtype=torch.float32, device=dev)
        self.step_count = 0
        self._norm_scratch = torch.empty(ops.grad_norm_scratch_bytes(arena.numel) // 4,
                                         dtype=torch.float32, device=dev)
        self._clip_out = torch.ones(2, dtype=torch.float32, device=dev)  # [total_norm, coef]
        self._clip_armed = False
        ids = {id(p) for g in self.param_groups for p in g["params"]}
        self._names = {}
        for name, slot in arena.slots.items():
            self._names[name] = slot
        if len(ids) != len(arena.slots):
            raise NsbError("FusedAdamW must be given every parameter of the arena exactly once")

    # -- clip_grad_norm_ replacement -----------------------------------------------------------
    def clip_grad_norm_(self, max_norm: float) -> torch.Tensor:
        """Computes the global L2 norm of the gradient arena and arms the clip coefficient
        min(1, max_norm / (norm + 1e-6)) for the next step().  Returns the norm (device scalar)."""
        ops.grad_clip_coef(self.arena.grads, max_norm, self._norm_scratch, self._clip_out)
        self._clip_armed = True
        return self._clip_out[0]

    @torch.no_grad()
    def step(self, closure=None):
        if closure is not None:
            raise NsbError("FusedAdamW.step does not support closures")
        g0, g1 = self.param_groups[0], self.param_groups[-1]
        for g in self.param_groups:
            if g["lr"] != g0["lr"] or tuple(g["betas"]) != tuple(g0["betas"]) or g["eps"] != g0["eps"]:
                raise NsbError("FusedAdamW: all groups must share lr/betas/eps (only weight_decay "
                               "may differ, as in nanoGPT's configure_optimizers)")
        wds = {g["weight_decay"] for g in self.param_groups if g["weight_decay"] != 0.0}
        if len(wds) > 1:
            raise NsbError("FusedAdamW: one non-zero weight_decay value is supported")
        wd = wds.pop() if wds else 0.0
        lr = float(g0["lr"])
        b1, b2 = g0["betas"]
        self.step_count += 1
        t = self.step_count
        bc1 = 1.0 - b1 ** t
        bc2 = 1.0 - b2 ** t
        a = self.arena
        ops.adamw_step(a.params, a.grads, self.exp_avg, self.exp_avg_sq, a.lowp, a.chunk_decay,
                       float(b1), float(b2), float(g0["eps"]), 1.0 - lr * wd, lr / bc1, math.sqrt(bc2),
                       self._clip_out[1:] if self._clip_armed else None)
        self._clip_armed = False
        if a.lowp is not None:
            a.lowp_version = a.params._version  # the kernel refreshed the bf16 shadow
        return None

    def zero_grad(self, set_to_none: bool = True):
        # The gradient arena stays allocated (p.grad are views of it): zero it in one shot.
        self.arena.grads.zero_()

    # -- checkpoint format compatible with torch.optim.AdamW (per-parameter state) ------------
    def state_dict(self):
        sd = super().state_dict()
        order = [p for g in self.param_groups for p in g["params"]]
        by_ptr = {self.arena.p(n).data_ptr(): n for n in self.arena.slots}
        state = {}
        for i, p in enumerate(order):
            n = by_ptr[p.data_ptr()]
            state[i] = {"step": torch.tensor(float(self.step_count)),
                        "exp_avg": self.arena.view(self.exp_avg, n).clone(),
                        "exp_avg_sq": self.arena.view(self.exp_avg_sq, n).clone()}
        sd["state"] = state
        return sd

    def load_state_dict(self, state_dict):
        order = [p for g in self.param_groups for p in g["params"]]
        by_ptr = {self.arena.p(n).data_ptr(): n for n in self.arena.slots}
        st = state_dict["state"]
        for i, p in enumerate(order):
            if i not in st:
                continue
            n = by_ptr[p.data_ptr()]
            self.arena.view(self.exp_avg, n).copy_(st[i]["exp_avg"])
            self.arena.view(self.exp_avg_sq, n).copy_(st[i]["exp_avg_sq"])
            self.step_count = int(float(st[i]["step"]))
        for g, sg in zip(self.param_groups, state_dict["param_groups"]):
            for k in ("lr", "betas", "eps", "weight_decay"):
                if k in sg:
                    g[k] = sg[k]


def clip_grad_norm_(optimizer: FusedAdamW, max_norm: float) -> torch.Tensor:
    """train.py-facing helper mirroring torch.nn.utils.clip_grad_norm_'s role in the loop."""
    return optimizer.clip_grad_norm_(max_norm)
