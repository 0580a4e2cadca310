"""Loss kernels on the bench shape: two separate passes vs the one-pass kernel (CUDA events, L2 flushed)."""
import torch
from nanosandbox_b200 import ops
M, V = 12288, 50304
logits = torch.randn(M, V, device="cuda").to(torch.bfloat16)
tg = torch.randint(0, V, (M,), device="cuda")
lse, rl, st = torch.empty(M, device="cuda"), torch.empty(M, device="cuda"), torch.empty(2, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def t(fn, n=5):
    ts = []
    for _ in range(n):
        flush.zero_(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b) * 1e3)
    return sorted(ts)[len(ts) // 2]
gb = M * V * 2 / 1e9
f = t(lambda: ops.cross_entropy_fwd(logits, tg, lse, rl, st, V))
b = t(lambda: ops.cross_entropy_bwd(logits, tg, lse, st, None, V))
logits.normal_()
u = t(lambda: (logits.normal_(), ops.cross_entropy_fwd_bwd(logits, tg, lse, rl, st, V))) - t(lambda: logits.normal_())
print(f"cross-entropy fwd {f:.1f} us ({gb / f * 1e3:.0f} GB/s)  bwd {b:.1f} us ({2 * gb / b * 1e3:.0f} GB/s)  sum {f + b:.1f} us | one pass {u:.1f} us ({2 * gb / u * 1e3:.0f} GB/s)")
