"""Runs the attention kernels a few times on the cfg4 shape (for ncu captures)."""
import sys
import torch
from nanosandbox_b200 import ops
B, T, H, D = 12, 1024, 12, 64
C = H * D
qkv = torch.randn(B, T, 3 * C, device="cuda").to(torch.bfloat16)
y = torch.empty(B, T, C, device="cuda", dtype=torch.bfloat16)
dy = torch.randn_like(y)
lse = torch.empty(B, H, T, device="cuda")
dqkv = torch.empty_like(qkv)
delta = torch.empty(B, H, T, device="cuda")
dq_acc = torch.zeros(B * T, C, device="cuda")
for _ in range(3):
    ops.attention_fwd(qkv, y, lse, B, T, H, D)
    if "--bwd" in sys.argv:
        ops.attention_bwd(qkv, y, dy, lse, dqkv, delta, B, T, H, D, dq_acc)
torch.cuda.synchronize()
print("ok")
