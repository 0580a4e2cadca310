"""attention backward: staged dQ (TMA reduce-add) vs direct vector reductions; CUDA events, back-to-back."""
import torch
from nanosandbox_b200 import _lib, ops
lib = _lib.load()
B, T, H, D = 12, 1024, 12, 64
C = H * D
qkv = torch.randn(B, T, 3 * C, device="cuda").to(torch.bfloat16)
y = torch.empty(B, T, C, device="cuda", dtype=torch.bfloat16); dy = torch.randn_like(y)
lse = torch.empty(B, H, T, device="cuda"); dqkv = torch.empty_like(qkv)
delta = torch.empty(B, H, T, device="cuda"); dq_acc = torch.zeros(B * T, C, device="cuda")
ops.attention_fwd(qkv, y, lse, B, T, H, D)
outs = {}
for mode in (3, 4, 3, 4):
    lib.nsb_attention_debug_config(mode)
    for _ in range(3): ops.attention_bwd(qkv, y, dy, lse, dqkv, delta, B, T, H, D, dq_acc)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(20): ops.attention_bwd(qkv, y, dy, lse, dqkv, delta, B, T, H, D, dq_acc)
    e1.record(); torch.cuda.synchronize()
    print("mode", mode, "(staged)" if mode == 3 else "(direct)", f"{e0.elapsed_time(e1) / 20 * 1e3:.1f} us per backward (3 kernels)")
    outs[mode] = dqkv.float().clone()
print("max |dq staged - dq direct|", (outs[3] - outs[4]).abs().max().item(), "ref max", outs[3].abs().max().item())
