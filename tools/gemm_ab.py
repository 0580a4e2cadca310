import sys, torch
from nanosandbox_b200 import _lib, ops
sys.path.insert(0, "tools")
from bench_kernels import timeit, DEV
lib = _lib.load(); bf = torch.bfloat16; E = _lib
Mt, Cc = 12288, 768
def run(name, M, N, K, a_k, b_k, epi, acc, dbgs):
    A = torch.randn((M, K) if a_k else (K, M), device=DEV).to(bf); B = torch.randn((N, K) if b_k else (K, N), device=DEV).to(bf)
    cdt = torch.float32 if epi in (E.EPI_ACCUM, E.EPI_RESIDUAL) else bf
    Cout = torch.zeros(M, N, device=DEV, dtype=cdt)
    aux = torch.randn(M, N, device=DEV) if epi == E.EPI_RESIDUAL else None
    out = []
    for d in dbgs:
        lib.nsb_gemm_debug_config(0, 0, d)
        t = min(timeit(lambda: ops.gemm(M, N, K, A, A.stride(0), a_k, B, B.stride(0), b_k, Cout, N, epi, aux=aux, accumulate=acc)) for _ in range(3))
        lib.nsb_gemm_debug_config(0, 0, 0)
        out.append(f"dbg{d}: {t*1e3:6.1f} us")
    print(f"{name:22s}", "  ".join(out), flush=True)
D = (8, 9, 0)   # m-fastest, n-fastest, default heuristic
V = 50304
run("lm_head dgrad", Mt, Cc, V, True, False, E.EPI_STORE, False, D)
run("lm_head wgrad", V, Cc, Mt, False, False, E.EPI_ACCUM, True, D)
run("lm_head fwd", Mt, V, Cc, True, True, E.EPI_STORE, False, D)
run("fc fwd store", Mt, 4*Cc, Cc, True, True, E.EPI_STORE, False, D)
run("fc wgrad", 4*Cc, Cc, Mt, False, False, E.EPI_ACCUM, True, D)
run("mlp proj res", Mt, Cc, 4*Cc, True, True, E.EPI_RESIDUAL, False, D)
