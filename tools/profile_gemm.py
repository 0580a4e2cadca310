"""Runs selected GEMM shapes a few times (for ncu captures).  usage: profile_gemm.py <case>"""
import sys
import torch
from nanosandbox_b200 import _lib, ops
E = _lib
bf = torch.bfloat16
Mt, Cc = 12288, 768
case = sys.argv[1] if len(sys.argv) > 1 else "res"
def mk(M, N, K, a_k, b_k, epi, acc=False):
    A = torch.randn((M, K) if a_k else (K, M), device="cuda").to(bf)
    B = torch.randn((N, K) if b_k else (K, N), device="cuda").to(bf)
    cdt = torch.float32 if epi in (E.EPI_ACCUM, E.EPI_RESIDUAL) else bf
    Cout = torch.zeros(M, N, device="cuda", dtype=cdt)
    aux = torch.randn(M, N, device="cuda", dtype=torch.float32 if epi == E.EPI_RESIDUAL else bf) if epi in (E.EPI_RESIDUAL, E.EPI_DGELU) else None
    C2 = torch.empty(M, N, device="cuda", dtype=bf) if epi == E.EPI_GELU else None
    return lambda: ops.gemm(M, N, K, A, A.stride(0), a_k, B, B.stride(0), b_k, Cout, N, epi, C2=C2, aux=aux, accumulate=acc)
cases = {
    "res": mk(Mt, Cc, Cc, True, True, E.EPI_RESIDUAL),
    "qkv": mk(Mt, 3 * Cc, Cc, True, True, E.EPI_STORE),
    "gelu": mk(Mt, 4 * Cc, Cc, True, True, E.EPI_GELU),
    "lmhead_dgrad": mk(Mt, Cc, 50304, True, False, E.EPI_STORE),
    "fc_wgrad": mk(4 * Cc, Cc, Mt, False, False, E.EPI_ACCUM, True),
}
f = cases[case]
for _ in range(3):
    f()
torch.cuda.synchronize()
print("ok")
