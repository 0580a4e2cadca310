"""Generates tests/golden/*.npz from the oracle (run here, on CPU; the GPU box only reads them).

    python tools/make_golden.py

1. tiny_step.npz  : one forward/backward of a 2-layer model (bias=True): inputs, logits, loss,
                    every parameter gradient -- known-answer vectors for the CUDA path.
2. curve_cfg1.npz : 200 optimizer steps of BASELINE.json configs[0] (4L/4H/128, block 64, batch 12,
                    vocab 65, fp32) on the seeded Markov stream: per-step loss, the first batches
                    (data-sampling bit-exactness) and the final parameters' checksum.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.model import GPT, GPTConfig  # noqa: E402
from oracle.train_ref import RefTrainer, make_markov_stream  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
os.makedirs(OUT, exist_ok=True)
torch.set_num_threads(1)  # fixed summation order

# ---- 1. tiny step
kw = dict(n_layer=2, n_head=2, n_embd=64, block_size=32, vocab_size=65, bias=True, dropout=0.0)
torch.manual_seed(1337)
m = GPT(GPTConfig(**kw), verbose=False)
with torch.no_grad():
    for p in m.parameters():
        if p.dim() == 1:
            p.add_(torch.randn_like(p) * 0.05)
g = torch.Generator().manual_seed(99)
idx = torch.randint(0, 65, (3, 32), generator=g)
tgt = torch.randint(0, 65, (3, 32), generator=g)
tgt[1, 7] = -1
logits, loss = m(idx, tgt)
loss.backward()
blob = {"idx": idx.numpy(), "tgt": tgt.numpy(), "logits": logits.detach().numpy(), "loss": np.float32(loss.item())}
for n, p in m.named_parameters():
    blob["param/" + n] = p.detach().numpy()
    blob["grad/" + n] = p.grad.numpy()
np.savez_compressed(os.path.join(OUT, "tiny_step.npz"), **blob)
print("tiny_step: loss", loss.item())

# ---- 2. 200-step curve, cfg1
cfg = GPTConfig(n_layer=4, n_head=4, n_embd=128, block_size=64, vocab_size=65, bias=False, dropout=0.0)
data = make_markov_stream(65, 1 << 16, seed=0)
tr = RefTrainer(cfg, data, batch_size=12, grad_accum=1, learning_rate=1e-3, weight_decay=0.1,
                betas=(0.9, 0.99), grad_clip=1.0, decay_lr=True, warmup_iters=10, lr_decay_iters=200,
                min_lr=1e-4, seed=1337)
first_x = [tr.X.numpy().copy()]
losses = []
for i in range(200):
    losses.append(tr.step())
    if i < 3:
        first_x.append(tr.X.numpy().copy())
chk = float(sum(p.detach().double().sum() for p in tr.model.parameters()))
np.savez_compressed(os.path.join(OUT, "curve_cfg1.npz"), losses=np.array(losses, dtype=np.float64),
                    first_x=np.stack(first_x), data_head=data[:4096], param_sum=np.float64(chk))
print("curve: first", losses[0], "last", losses[-1], "param_sum", chk)
