"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py) -- CPU restatement of nanoGPT ``train.py``'s
training step: synthetic token streams, ``get_batch``, ``get_lr``, grad accumulation, clip,
AdamW, stock DistributedDataParallel.

Reference call sites: /root/reference/notebooks/colab_nanoGPT_companion.ipynb:70-79 (single
process), :107-116 (``torchrun --standalone --nproc_per_node=2``); launch contract
/root/reference/README.md:6-8, 101-102.  Arithmetic: torch/optim/adam.py:347-547 (AdamW),
torch/nn/utils/clip_grad.py:50-118,165-174 (clip), torch/nn/parallel/distributed.py (DDP).
"""
from __future__ import annotations

import math
import os
import pickle
from contextlib import nullcontext

import numpy as np
import torch

from .model import GPT, GPTConfig

STREAM_LEN = 1 << 24


def make_uniform_stream(vocab_size: int, n: int = STREAM_LEN, seed: int = 0) -> np.ndarray:
    """Uniform i.i.d. token stream (throughput runs; loss floor ln V).  SURVEY.md 8d."""
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, vocab_size, (n,), generator=g).numpy().astype(np.uint16)


def make_markov_stream(vocab_size: int, n: int = 1 << 20, seed: int = 0,
                       branching: int = 4) -> np.ndarray:
    """Learnable stream: order-1 Markov chain, each symbol has `branching` successors with
    fixed probabilities, so the loss can fall well below ln V (loss-curve parity)."""
    rng = np.random.RandomState(seed)
    succ = rng.randint(0, vocab_size, size=(vocab_size, branching))
    probs = np.array([0.6, 0.2, 0.15, 0.05][:branching], dtype=np.float64)
    probs = probs / probs.sum()
    choice = rng.choice(branching, size=n, p=probs)
    out = np.empty(n, dtype=np.uint16)
    s = 0
    for i in range(n):
        s = succ[s, choice[i]]
        out[i] = s
    return out


def write_dataset(data_dir: str, vocab_size: int, kind: str = "uniform",
                  n: int = STREAM_LEN, seed: int = 0, write_meta: bool | None = None) -> None:
    """train.bin / val.bin (uint16) + meta.pkl, the files nanoGPT's prepare.py leaves behind
    (/root/reference/notebooks/colab_nanoGPT_companion.ipynb:50-57)."""
    os.makedirs(data_dir, exist_ok=True)
    mk = make_uniform_stream if kind == "uniform" else make_markov_stream
    mk(vocab_size, n, seed).tofile(os.path.join(data_dir, "train.bin"))
    mk(vocab_size, max(n // 8, 4096), seed + 1).tofile(os.path.join(data_dir, "val.bin"))
    if write_meta if write_meta is not None else vocab_size < 50000:
        with open(os.path.join(data_dir, "meta.pkl"), "wb") as f:
            pickle.dump({"vocab_size": vocab_size}, f)


def get_batch(data: np.ndarray, batch_size: int, block_size: int, device="cpu"):
    """Draws from the GLOBAL CPU generator, exactly like nanoGPT (a1 in SURVEY.md 8a)."""
    ix = torch.randint(len(data) - block_size, (batch_size,))
    x = torch.stack([torch.from_numpy(data[i:i + block_size].astype(np.int64)) for i in ix])
    y = torch.stack([torch.from_numpy(data[i + 1:i + 1 + block_size].astype(np.int64)) for i in ix])
    return x.to(device), y.to(device)


def get_lr(it: int, learning_rate: float, warmup_iters: int, lr_decay_iters: int,
           min_lr: float) -> float:
    """Linear warm-up then cosine decay.  Pinned form: lr*(it+1)/(warmup+1) (current upstream;
    SURVEY.md a16 notes older revisions used lr*it/warmup)."""
    if it < warmup_iters:
        return learning_rate * (it + 1) / (warmup_iters + 1)
    if it > lr_decay_iters:
        return min_lr
    ratio = (it - warmup_iters) / (lr_decay_iters - warmup_iters)
    return min_lr + 0.5 * (1.0 + math.cos(math.pi * ratio)) * (learning_rate - min_lr)


class RefTrainer:
    """The hot loop of train.py for one rank, on stock torch."""

    def __init__(self, cfg: GPTConfig, data: np.ndarray, batch_size: int, grad_accum: int = 1,
                 learning_rate=6e-4, weight_decay=0.1, betas=(0.9, 0.95), grad_clip=1.0,
                 warmup_iters=0, lr_decay_iters=1000, min_lr=6e-5, decay_lr=False,
                 seed=1337, device="cpu", dtype="float32", ddp=False, verbose=False):
        torch.manual_seed(seed)
        self.cfg, self.data, self.B, self.accum = cfg, data, batch_size, grad_accum
        self.device = device
        self.model = GPT(cfg, verbose=verbose).to(device)
        dev_type = "cuda" if "cuda" in str(device) else "cpu"
        self.opt = self.model.configure_optimizers(weight_decay, learning_rate, betas, dev_type,
                                                   verbose=verbose)
        pt = {"float32": torch.float32, "bfloat16": torch.bfloat16}[dtype]
        self.ctx = (nullcontext() if dtype == "float32"
                    else torch.amp.autocast(device_type=dev_type, dtype=pt))
        self.lr_args = (learning_rate, warmup_iters, lr_decay_iters, min_lr)
        self.decay_lr, self.grad_clip, self.lr = decay_lr, grad_clip, learning_rate
        self.wrapped = self.model
        if ddp:
            from torch.nn.parallel import DistributedDataParallel as DDP
            self.wrapped = DDP(self.model, device_ids=None if dev_type == "cpu" else [device])
        self.ddp = ddp
        self.iter_num = 0
        self.X, self.Y = get_batch(data, batch_size, cfg.block_size, device)

    def step(self) -> float:
        lr = get_lr(self.iter_num, *self.lr_args) if self.decay_lr else self.lr
        for g in self.opt.param_groups:
            g["lr"] = lr
        total = 0.0
        for micro in range(self.accum):
            if self.ddp:
                self.wrapped.require_backward_grad_sync = (micro == self.accum - 1)
            with self.ctx:
                _, loss = self.wrapped(self.X, self.Y)
                loss = loss / self.accum
            self.X, self.Y = get_batch(self.data, self.B, self.cfg.block_size, self.device)
            loss.backward()
            total += float(loss.detach())
        if self.grad_clip != 0.0:
            torch.nn.utils.clip_grad_norm_(self.model.parameters(), self.grad_clip)
        self.opt.step()
        self.opt.zero_grad(set_to_none=True)
        self.iter_num += 1
        return total
