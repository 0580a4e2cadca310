"""Char GPT (6L/6H/384, block 256, batch 64) optimizer step with and without dropout (GPU box)."""
import time
import torch
from nanosandbox_b200.model import GPT, GPTConfig
for p in (0.0, 0.2):
    torch.manual_seed(0)
    m = GPT(GPTConfig(n_layer=6, n_head=6, n_embd=384, block_size=256, vocab_size=65, bias=False, dropout=p), verbose=False).cuda()
    opt = m.configure_optimizers(0.1, 1e-3, (0.9, 0.99), "cuda", verbose=False)
    x = torch.randint(0, 65, (64, 256), device="cuda")
    def step():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            _, loss = m(x, x)
        loss.backward(); opt.clip_grad_norm_(1.0); opt.step(); opt.zero_grad()
    for _ in range(5): step()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(30): step()
    torch.cuda.synchronize(); print(f"char step, dropout={p}: {(time.perf_counter()-t0)/30*1e3:.2f} ms")
