"""In-kernel traces of one tcgen05 GEMM launch: clock64 stamps of CTA 0 (MMA warp per tile, epilogue
warp 0 per chunk) and wall-clock (globaltimer) stamps of every CTA's life.
usage: gemm_timing.py M N K [store|gelu|res|accum] [a_kmajor b_kmajor]"""
import sys
import torch
from nanosandbox_b200 import _lib, ops

lib = _lib.load(); bf = torch.bfloat16; E = _lib
M, N, K = (int(x) for x in sys.argv[1:4])
kind = sys.argv[4] if len(sys.argv) > 4 else "store"
a_k = int(sys.argv[5]) if len(sys.argv) > 5 else 1
b_k = int(sys.argv[6]) if len(sys.argv) > 6 else 1
epi = {"store": E.EPI_STORE, "gelu": E.EPI_GELU, "res": E.EPI_RESIDUAL, "accum": E.EPI_ACCUM, "dgelu": E.EPI_DGELU}[kind]
A = (torch.randn(M, K, device="cuda") if a_k else torch.randn(K, M, device="cuda")).to(bf)
B = (torch.randn(N, K, device="cuda") if b_k else torch.randn(K, N, device="cuda")).to(bf)
cdt = torch.float32 if epi in (E.EPI_RESIDUAL, E.EPI_ACCUM) else bf
C = torch.zeros(M, N, device="cuda", dtype=cdt); C2 = torch.zeros(M, N, device="cuda", dtype=bf)
aux = torch.randn(M, N, device="cuda") if epi == E.EPI_RESIDUAL else (torch.randn(M, N, device="cuda").to(bf) if epi == E.EPI_DGELU else None)
dbg = torch.zeros(8192, dtype=torch.int64, device="cuda")
f = lambda: ops.gemm(M, N, K, A, K if a_k else M, bool(a_k), B, K if b_k else N, bool(b_k), C, N, epi,
                     C2=C2 if epi == E.EPI_GELU else None, aux=aux)
import os
if os.environ.get("GEMM_BN"): lib.nsb_gemm_debug_config(int(os.environ["GEMM_BN"]), 0, int(os.environ.get("GEMM_MODE", "0")))
elif os.environ.get("GEMM_MODE"): lib.nsb_gemm_debug_config(0, 0, int(os.environ["GEMM_MODE"]))
if os.environ.get("GEMM_PIPE"): lib.nsb_gemm_debug_pipeline(*[int(x) for x in os.environ["GEMM_PIPE"].split(",")])
f(); f()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for rep in (5, 50):
    torch.cuda.synchronize(); e0.record()
    for _ in range(rep): f()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / rep
    print(f"{M}x{N}x{K} {kind} {rep} back-to-back: {us:.1f} us  {2 * M * N * K / us / 1e6:.0f} TF/s")
torch.cuda.synchronize()
lib.nsb_gemm_debug_timing(dbg.data_ptr()); f(); torch.cuda.synchronize(); lib.nsb_gemm_debug_timing(None)
full = dbg.cpu().tolist()
m = full[1024:1024 + 240]
print("MMA warp per tile: [start, acc free, all k-blocks issued] rel. clocks; wait-for-operands cycles")
for i in range(0, 60):
    r = m[4 * i:4 * i + 4]
    if r[0] == 0: break
    if i < 4:
        print(" tile", i, [x - m[0] for x in r[:3]], "tempty wait", r[1] - r[0], "mainloop", r[2] - r[1], "of which waiting for TMA", r[3])
g = [x for x in full[2048:2048 + 60] if x != 0]
if len(g) > 2:
    nt = len(g)
    print(f"effective SM clock over {nt - 1} tiles: {(m[4 * (nt - 1)] - m[0]) / (g[-1] - g[0]):.3f} GHz; tile period {(g[-1] - g[0]) / (nt - 1) / 1e3:.2f} us")
ct = [(i, full[4096 + 8 * i:4096 + 8 * i + 8]) for i in range(148) if full[4096 + 8 * i] != 0]
if ct:
    t0 = min(c[0] for _, c in ct)
    names = ["entry", "setup", "pdl", "exit", "1st operands", "mma issued", "epi staged", "stores read"]
    for k in (0, 1, 2, 4, 5, 6, 7, 3):
        vals = [c[k] - t0 for _, c in ct if c[k] != 0]
        if vals:
            vals.sort()
            print(f"  {names[k]:13s} min {vals[0]:7d}  median {vals[len(vals) // 2]:7d}  max {vals[-1]:7d} ns   (n={len(vals)})")
    late = sorted(ct, key=lambda ic: -ic[1][3])[:6]
    print("  latest CTAs:", [(i, [x - t0 if x else 0 for x in c]) for i, c in late])
    early = sorted(ct, key=lambda ic: ic[1][3])[:3]
    print("  earliest CTAs:", [(i, [x - t0 if x else 0 for x in c]) for i, c in early])
d = [x for x in full[:1000] if x != 0]
if not d:
    sys.exit(0)        # this epilogue path records no per-chunk stamps
t0 = d[0]
rows = [d[i:i + 4] for i in range(0, len(d) - 3, 4)]
for i, r in enumerate(rows[:int(sys.argv[7]) if len(sys.argv) > 7 else 10]):
    print(i, [x - t0 for x in r], "ld+acquire", r[1] - r[0], "math+stage", r[2] - r[1], "publish", r[3] - r[2],
          "gap-to-next", (rows[i + 1][0] - r[3]) if i + 1 < len(rows) else None)
