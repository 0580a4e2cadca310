"""Attention backward at the cfg4 shape: persistent walk (mode 17, the default) against one CTA per item (16)."""
import sys
import torch
from nanosandbox_b200 import _lib, ops
lib = _lib.load()
B, T, H, D = 12, 1024, 12, 64
if len(sys.argv) > 3:
    B, T, H = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
C = H * D
torch.manual_seed(0)
qkv = torch.randn(B, T, 3 * C, device="cuda").to(torch.bfloat16)
y = torch.empty(B, T, C, device="cuda", dtype=torch.bfloat16); dy = torch.randn_like(y)
lse = torch.empty(B, H, T, device="cuda"); delta = torch.empty(B, H, T, device="cuda")
dq_acc = torch.zeros(B * T, C, device="cuda")
ops.attention_fwd(qkv, y, lse, B, T, H, D)
def t(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) * 1e3 / n
outs = {}
for mode in (16, 17, 16, 17):
    lib.nsb_attention_debug_config(mode)
    dqkv = torch.zeros_like(qkv)
    us = t(lambda: ops.attention_bwd(qkv, y, dy, lse, dqkv, delta, B, T, H, D, dq_acc))
    outs[mode] = dqkv.float()
    print(f"mode {mode}: bwd (3 kernels) {us:.1f} us")
d = (outs[16] - outs[17]).abs().max().item()
print("max |difference| between the two forms:", d, " max |value|:", outs[16].abs().max().item())
lib.nsb_attention_debug_config(17)
