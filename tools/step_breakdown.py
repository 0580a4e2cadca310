"""In-situ time per kernel family inside one real optimizer step (CUDA events around every C-ABI
call, same stream, warm caches, back-to-back launches) for the bench workload."""
import collections
import json
import sys

import torch

import bench
from nanosandbox_b200 import ops
import nanosandbox_b200.model as model_mod
import nanosandbox_b200.optim as optim_mod
from nanosandbox_b200.model import GPT, GPTConfig

dev = torch.device("cuda", 0)
torch.manual_seed(1337)
model = GPT(GPTConfig(**bench.CFG), verbose=False).to(dev)
opt = model.configure_optimizers(0.1, 6e-4, (0.9, 0.95), "cuda", verbose=False)
T, V = bench.CFG["block_size"], bench.CFG["vocab_size"]
micro = int(sys.argv[1]) if len(sys.argv) > 1 else 4
batches = [(torch.randint(0, V, (bench.BATCH, T), device=dev), torch.randint(0, V, (bench.BATCH, T), device=dev)) for _ in range(micro)]


def step():
    for X, Y in batches:
        with torch.autocast("cuda", dtype=torch.bfloat16):
            _, loss = model(X, Y)
        (loss / micro).backward()
    opt.clip_grad_norm_(1.0)
    opt.step()
    opt.zero_grad(set_to_none=True)


for _ in range(3):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); step(); e1.record(); torch.cuda.synchronize()
plain_ms = e0.elapsed_time(e1)

events = []
names = ["embedding_fwd", "embedding_bwd", "layernorm_fwd", "layernorm_bwd", "cross_entropy_fwd", "cross_entropy_bwd",
         "grad_clip_coef", "adamw_step", "cast_f32_to_bf16", "gemm", "attention_fwd", "attention_bwd", "gelu_fwd", "gelu_bwd"]
orig = {n: getattr(ops, n) for n in names}


def wrap(n):
    f = orig[n]

    def g(*a, **k):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); f(*a, **k); e.record()
        tag = n
        if n == "gemm":
            M, N, K = a[0], a[1], a[2]
            epi = a[11] if len(a) > 11 else k.get("epilogue", 0)
            tag = f"gemm {M}x{N}x{K} epi{epi}"
        events.append((tag, s, e))
    return g


for n in names:
    setattr(ops, n, wrap(n))
step()
torch.cuda.synchronize()
for n in names:
    setattr(ops, n, orig[n])
agg = collections.defaultdict(lambda: [0, 0.0])
for tag, s, e in events:
    agg[tag][0] += 1
    agg[tag][1] += s.elapsed_time(e)
tot = sum(v[1] for v in agg.values())
print(f"plain step {plain_ms:.2f} ms for {micro} micro-steps ({plain_ms / micro:.2f} ms each); sum of timed calls {tot:.2f} ms")
fam = collections.defaultdict(float)
for tag, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{tag:34s} n={n:4d} total {ms:8.3f} ms  avg {ms / n * 1e3:8.1f} us  {ms / tot * 100:5.1f}%")
    fam[tag.split()[0]] += ms
print("families:", {k: round(v / micro, 3) for k, v in sorted(fam.items(), key=lambda kv: -kv[1])}, "ms per micro-step")
