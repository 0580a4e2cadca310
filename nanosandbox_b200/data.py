"""Synthetic token streams in nanoGPT's on-disk format (train.bin / val.bin of uint16 + meta.pkl).

The reference's jobs run nanoGPT's ``prepare.py`` in a download Job
(/root/reference/README.md:48-53; notebook :50-57).  There is no network here, so throughput and
parity runs use seeded synthetic streams of the same format: a uniform i.i.d.
stream (loss floor ln V) and a learnable order-1 Markov stream for loss-curve checks.

    python -m nanosandbox_b200.data --out data/shakespeare_char --vocab 65 [--kind markov]
"""
from __future__ import annotations

import argparse
import os
import pickle

import numpy as np
import torch

STREAM_LEN = 1 << 24


def uniform_stream(vocab_size: int, n: int = STREAM_LEN, seed: int = 0) -> np.ndarray:
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, vocab_size, (n,), generator=g).numpy().astype(np.uint16)


def markov_stream(vocab_size: int, n: int = 1 << 20, seed: int = 0, branching: int = 4,
                  chain_seed: int | None = None) -> np.ndarray:
    """Order-1 Markov chain over the vocabulary.  chain_seed=None: transition table and walk from one generator
    seeded with `seed` (the form the golden curves were made with); otherwise the table comes from `chain_seed`
    and the walk from `seed`, so that a validation stream can follow the SAME chain as the training stream."""
    rng = np.random.RandomState(seed)
    if chain_seed is None:
        succ = rng.randint(0, vocab_size, size=(vocab_size, branching))
    else:
        succ = np.random.RandomState(chain_seed).randint(0, vocab_size, size=(vocab_size, branching))
    probs = np.array([0.6, 0.2, 0.15, 0.05][:branching], dtype=np.float64)
    probs /= probs.sum()
    choice = rng.choice(branching, size=n, p=probs)
    out = np.empty(n, dtype=np.uint16)
    s = 0
    for i in range(n):
        s = succ[s, choice[i]]
        out[i] = s
    return out


def write_dataset(data_dir: str, vocab_size: int, kind: str = "uniform", n: int | None = None,
                  seed: int = 0) -> None:
    os.makedirs(data_dir, exist_ok=True)
    if kind == "uniform":
        n = n or STREAM_LEN
        tr, va = uniform_stream(vocab_size, n, seed), uniform_stream(vocab_size, max(n // 8, 4096), seed + 1)
    elif kind == "markov":
        n = n or (1 << 20)
        tr = markov_stream(vocab_size, n, seed)
        va = markov_stream(vocab_size, max(n // 8, 4096), seed + 1, chain_seed=seed)   # same chain, another walk
    else:
        raise ValueError(f"unknown synthetic kind {kind!r}")
    # write under temporary names and rename into place (val.bin, then meta.pkl, train.bin last: readers
    # test for train.bin), so that a concurrent reader never maps a partly written file
    tag = f".tmp{os.getpid()}"
    final = []
    va.tofile(os.path.join(data_dir, "val.bin" + tag))
    final.append("val.bin")
    if vocab_size < 50000:  # char-level datasets carry their vocab in meta.pkl; GPT-2 BPE ones do not
        with open(os.path.join(data_dir, "meta.pkl" + tag), "wb") as f:
            pickle.dump({"vocab_size": vocab_size}, f)
        final.append("meta.pkl")
    tr.tofile(os.path.join(data_dir, "train.bin" + tag))
    final.append("train.bin")
    for name in final:
        os.replace(os.path.join(data_dir, name + tag), os.path.join(data_dir, name))


def get_batch(data: np.ndarray, batch_size: int, block_size: int, device, pin: bool = True):
    """nanoGPT's get_batch: indices from the GLOBAL CPU generator, int64 x/y, pinned async H2D."""
    ix = torch.randint(len(data) - block_size, (batch_size,))
    x = torch.stack([torch.from_numpy(data[i:i + block_size].astype(np.int64)) for i in ix])
    y = torch.stack([torch.from_numpy(data[i + 1:i + 1 + block_size].astype(np.int64)) for i in ix])
    if pin and str(device).startswith("cuda"):
        return x.pin_memory().to(device, non_blocking=True), y.pin_memory().to(device, non_blocking=True)
    return x.to(device), y.to(device)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True)
    ap.add_argument("--vocab", type=int, required=True)
    ap.add_argument("--kind", default="uniform", choices=["uniform", "markov"])
    ap.add_argument("--n", type=int, default=0)
    a = ap.parse_args()
    write_dataset(a.out, a.vocab, a.kind, a.n or None)
    print(f"wrote {a.out}/train.bin, val.bin")
