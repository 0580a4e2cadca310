"""One bf16 training forward/backward on the other BASELINE configs' shapes (char: 6L/6H/384, T=256; GPT-2 medium:
24L/16H/1024, T=1024) against the oracle model on the same GPU: loss and a few gradient tensors."""
import torch
from nanosandbox_b200.model import GPT, GPTConfig
from oracle.model import GPT as RefGPT, GPTConfig as RefConfig

def rel(a, b): return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-30)).item()

for name, kw, B in (("char", dict(n_layer=6, n_head=6, n_embd=384, block_size=256, vocab_size=65, bias=False, dropout=0.0), 64),
                    ("gpt2-medium", dict(n_layer=24, n_head=16, n_embd=1024, block_size=1024, vocab_size=50304, bias=False, dropout=0.0), 4)):
    torch.manual_seed(1337); ref = RefGPT(RefConfig(**kw), verbose=False).to("cuda")
    torch.manual_seed(1337); ours = GPT(GPTConfig(**kw), verbose=False).to("cuda")
    g = torch.Generator().manual_seed(0)
    T, V = kw["block_size"], kw["vocab_size"]
    idx = torch.randint(0, V, (B, T), generator=g).to("cuda"); tgt = torch.randint(0, V, (B, T), generator=g).to("cuda")
    with torch.autocast("cuda", dtype=torch.bfloat16):
        _, rl = ref(idx, tgt); _, ol = ours(idx, tgt)
    rl.backward(); ol.backward(); torch.cuda.synchronize()
    errs = {n: rel(q.grad, p.grad) for (n, p), (_, q) in zip(ref.named_parameters(), ours.named_parameters())}
    worst = max(errs, key=errs.get)
    print(f"{name}: loss ours {ol.item():.5f} oracle {rl.item():.5f}; worst gradient rel err {errs[worst]:.3e} ({worst}); median {sorted(errs.values())[len(errs)//2]:.3e}")
    del ref, ours
    torch.cuda.empty_cache()
