"""attention forward / backward timing on the bench shape (CUDA events, back-to-back launches)."""
import torch
from nanosandbox_b200 import _lib, ops
B, T, H, D = 12, 1024, 12, 64
C = H * D
qkv = torch.randn(B, T, 3 * C, device="cuda").to(torch.bfloat16)
y = torch.empty(B, T, C, device="cuda", dtype=torch.bfloat16); dy = torch.randn_like(y)
lse = torch.empty(B, H, T, device="cuda"); dqkv = torch.empty_like(qkv)
delta = torch.empty(B, H, T, device="cuda"); dq_acc = torch.zeros(B * T, C, device="cuda")
def t(fn, n=20):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
for _ in range(2):
    print(f"attention fwd {t(lambda: ops.attention_fwd(qkv, y, lse, B, T, H, D)):.1f} us   bwd (3 kernels) {t(lambda: ops.attention_bwd(qkv, y, dy, lse, dqkv, delta, B, T, H, D, dq_acc)):.1f} us")
lib = _lib.load()
outs = {}
for mode in (14, 15, 14, 15):
    lib.nsb_attention_debug_config(mode)
    us = t(lambda: ops.attention_fwd(qkv, y, lse, B, T, H, D))
    print("forward,", "4 softmax warps" if mode == 14 else "8 softmax warps", f"{us:.1f} us")
    outs[mode] = (y.float().clone(), lse.clone())
print("max |diff| y:", (outs[14][0] - outs[15][0]).abs().max().item(), " lse:", (outs[14][1] - outs[15][1]).abs().max().item(), " y max", outs[14][0].abs().max().item())
