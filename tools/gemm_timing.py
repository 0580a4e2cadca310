"""clock64 stamps of CTA 0's epilogue thread 0: per chunk [start, staged, drained, barrier]."""
import sys, torch
from nanosandbox_b200 import _lib, ops
lib = _lib.load(); bf = torch.bfloat16; E = _lib
M, N, K = 12288, 3072, int(sys.argv[1]) if len(sys.argv) > 1 else 768
epi = {"store": E.EPI_STORE, "gelu": E.EPI_GELU, "res": E.EPI_RESIDUAL}[sys.argv[2] if len(sys.argv) > 2 else "store"]
if epi == E.EPI_RESIDUAL: N = 768
A = torch.randn(M, K, device="cuda").to(bf); B = torch.randn(N, K, device="cuda").to(bf)
cdt = torch.float32 if epi == E.EPI_RESIDUAL else bf
C = torch.zeros(M, N, device="cuda", dtype=cdt); C2 = torch.zeros(M, N, device="cuda", dtype=bf)
aux = torch.randn(M, N, device="cuda") if epi == E.EPI_RESIDUAL else None
dbg = torch.zeros(4096, dtype=torch.int64, device="cuda")
f = lambda: ops.gemm(M, N, K, A, K, True, B, K, True, C, N, epi, C2=C2 if epi == E.EPI_GELU else None, aux=aux)
f(); f()
lib.nsb_gemm_debug_timing(dbg.data_ptr()); f(); torch.cuda.synchronize(); lib.nsb_gemm_debug_timing(None)
full = dbg.cpu().tolist()
m = full[1024:1024 + 240]
print("MMA warp per tile: [start, acc free, all k-blocks issued] rel. clocks; wait-for-operands cycles")
for i in range(0, 60):
    r = m[4 * i:4 * i + 4]
    if r[0] == 0: break
    print(" tile", i, [x - m[0] for x in r[:3]], "tempty wait", r[1] - r[0], "mainloop", r[2] - r[1], "of which waiting for TMA", r[3])
d = [x for x in full[:1000] if x != 0]
t0 = d[0]
rows = [d[i:i + 4] for i in range(0, len(d) - 3, 4)]
for i, r in enumerate(rows[:40]):
    print(i, [x - t0 for x in r], "tmem+math", r[1] - r[0], "drain-wait", r[2] - r[1], "barrier", r[3] - r[2],
          "gap-to-next", (rows[i + 1][0] - r[3]) if i + 1 < len(rows) else None)
