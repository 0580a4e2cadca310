"""LayerNorm backward on the bench shape (CUDA events, back-to-back launches, warm L2 as in the step)."""
import torch
from nanosandbox_b200 import ops
M, C = 12288, 768
x = torch.randn(M, C, device="cuda"); w = torch.ones(C, device="cuda")
yb = torch.empty(M, C, device="cuda", dtype=torch.bfloat16)
mean = torch.empty(M, device="cuda"); rstd = torch.empty(M, device="cuda")
dyb = torch.randn(M, C, device="cuda").to(torch.bfloat16)
dres = torch.randn(M, C, device="cuda"); dx = torch.empty_like(dres); dxl = torch.empty_like(dyb)
dw = torch.zeros(C, device="cuda")
scratch = torch.empty(ops.layernorm_bwd_scratch_bytes(M, C) // 4, device="cuda")
ops.layernorm_fwd(x, w, None, yb, mean, rstd)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def t(fn, n=20, cold=False):
    for _ in range(3): fn()
    ts = []
    for _ in range(n):
        if cold: flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    return sorted(ts)[len(ts) // 2]
f = lambda: ops.layernorm_bwd(dyb, x, w, mean, rstd, dres, dx, dxl, dw, None, True, scratch)
print(f"layernorm bwd (2 kernels): warm {t(f):.1f} us, L2 flushed {t(f, cold=True):.1f} us; fwd warm {t(lambda: ops.layernorm_fwd(x, w, None, yb, mean, rstd)):.1f} us")
