import torch
from nanosandbox_b200 import _lib, ops
lib = _lib.load(); bf = torch.bfloat16
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timeit(fn, iters=10):
    for _ in range(3): fn()
    ts = []
    for _ in range(iters):
        flush.zero_(); torch.cuda._sleep(200_000)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2] * 1e3
K = 12288
for M, N in ((2304, 768), (768, 768), (3072, 768), (768, 3072)):
    A = torch.randn(K, M, device="cuda").to(bf); B = torch.randn(K, N, device="cuda").to(bf)
    C = torch.zeros(M, N, device="cuda")
    for bn in (0, 128, 256):
        row = []
        for sp in (0, 1, 2, 3, 4, 6, 8):
            lib.nsb_gemm_debug_config(bn, sp, 0)
            t = timeit(lambda: ops.gemm(M, N, K, A, M, False, B, N, False, C, N, _lib.EPI_ACCUM, accumulate=True))
            row.append(f"{sp}:{t:.1f}")
        lib.nsb_gemm_debug_config(0, 0, 0)
        print(f"wgrad {M}x{N}x{K} bn={bn} splits:us  " + "  ".join(row), flush=True)
