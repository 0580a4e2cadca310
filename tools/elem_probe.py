"""One launch of each memory-bound kernel at the cfg4 shapes (for an ncu source-level capture)."""
import torch
from nanosandbox_b200 import ops
DEV = "cuda"; Mt, Cc, V = 12288, 768, 50304
x = torch.randn(Mt, Cc, device=DEV); w = torch.ones(Cc, device=DEV)
yb = torch.empty(Mt, Cc, device=DEV, dtype=torch.bfloat16); mean = torch.empty(Mt, device=DEV); rstd = torch.empty(Mt, device=DEV)
dyb = torch.randn(Mt, Cc, device=DEV).to(torch.bfloat16); dres = torch.randn(Mt, Cc, device=DEV); dx = torch.empty_like(dres); dxl = torch.empty_like(dyb)
dw = torch.zeros(Cc, device=DEV); scratch = torch.empty(ops.layernorm_bwd_scratch_bytes(Mt, Cc) // 4, device=DEV)
logits = torch.randn(Mt, V, device=DEV).to(torch.bfloat16); tg = torch.randint(0, V, (Mt,), device=DEV)
rl = torch.empty(Mt, device=DEV); ls = torch.empty(Mt, device=DEV); st = torch.empty(2, device=DEV)
hh = torch.randn(Mt, 4 * Cc, device=DEV).to(torch.bfloat16); aa = torch.empty_like(hh)
for _ in range(3):
    ops.layernorm_fwd(x, w, None, yb, mean, rstd)
    ops.layernorm_bwd(dyb, x, w, mean, rstd, dres, dx, dxl, dw, None, True, scratch)
    ops.gelu_fwd(hh, aa); ops.gelu_bwd(aa, hh)
    ops.cross_entropy_fwd_bwd(logits, tg, ls, rl, st, V)
torch.cuda.synchronize(); print("done")
