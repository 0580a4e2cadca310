"""lm_head forward GEMM (12288 x 50304 x 768, 1.2 GB of bf16 logits) with and without the evict_first store hint."""
import torch
from nanosandbox_b200 import _lib, ops
lib = _lib.load(); bf = torch.bfloat16
M, N, K = 12288, 50304, 768
A = torch.randn(M, K, device="cuda").to(bf); B = torch.randn(N, K, device="cuda").to(bf)
C = torch.empty(M, N, device="cuda", dtype=bf)
def t(n=20):
    for _ in range(3): ops.gemm(M, N, K, A, K, True, B, K, True, C, N)
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): ops.gemm(M, N, K, A, K, True, B, K, True, C, N)
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) * 1e3 / n
for rep in range(2):
    for on in (0, 1):
        lib.nsb_gemm_debug_store_hint(on)
        print(f"store hint {on}: {t():.1f} us", flush=True)
ref = (A[:256].float() @ B.float().t())
ops.gemm(M, N, K, A, K, True, B, K, True, C, N); torch.cuda.synchronize()
print("rel err first 256 rows", ((C[:256].float() - ref).norm() / ref.norm()).item())
