import torch
from nanosandbox_b200 import ops
M, N = 12288, 3072
h = torch.randn(M, N, device="cuda").to(torch.bfloat16); a = torch.empty_like(h)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def t(fn, n=7):
    ts = []
    for _ in range(n):
        flush.zero_(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    return sorted(ts)[len(ts) // 2]
gb = M * N * 2 / 1e9
f = t(lambda: ops.gelu_fwd(h, a)); b = t(lambda: ops.gelu_bwd(a, h))
print(f"gelu fwd {f:.1f} us ({2 * gb / f * 1e3:.0f} GB/s)  gelu bwd {b:.1f} us ({3 * gb / b * 1e3:.0f} GB/s)  [L2 flushed]")
