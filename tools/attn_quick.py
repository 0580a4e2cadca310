"""Attention forward / backward time at the cfg4 shape (B=12, T=1024, H=12, D=64), back-to-back launches."""
import torch
from nanosandbox_b200 import ops
B, T, H, D = 12, 1024, 12, 64
C = H * D
qkv = torch.randn(B, T, 3 * C, device="cuda").to(torch.bfloat16)
y = torch.empty(B, T, C, device="cuda", dtype=torch.bfloat16); dy = torch.randn_like(y)
lse = torch.empty(B, H, T, device="cuda"); dqkv = torch.empty_like(qkv); delta = torch.empty(B, H, T, device="cuda")
dq_acc = torch.zeros(B * T, C, device="cuda")
def t(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) * 1e3 / n
print(f"attention fwd {t(lambda: ops.attention_fwd(qkv, y, lse, B, T, H, D)):.1f} us   bwd (3 kernels) {t(lambda: ops.attention_bwd(qkv, y, dy, lse, dqkv, delta, B, T, H, D, dq_acc)):.1f} us")
