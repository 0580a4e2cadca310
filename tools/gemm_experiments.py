"""Tuning experiments for the tcgen05 GEMM (debug knobs): N tile, epilogue groups, split-K."""
import sys
import torch
from nanosandbox_b200 import _lib, ops
sys.path.insert(0, "tools")
from bench_kernels import timeit, DEV

lib = _lib.load()
bf = torch.bfloat16


def run(name, M, N, K, a_k, b_k, epi, accumulate=False, cfgs=()):
    A = torch.randn((M, K) if a_k else (K, M), device=DEV).to(bf)
    B = torch.randn((N, K) if b_k else (K, N), device=DEV).to(bf)
    cdt = torch.float32 if epi in (_lib.EPI_ACCUM, _lib.EPI_RESIDUAL) else bf
    Cout = torch.zeros(M, N, device=DEV, dtype=cdt)
    aux = torch.randn(M, N, device=DEV, dtype=torch.float32 if epi == _lib.EPI_RESIDUAL else bf) if epi in (_lib.EPI_RESIDUAL, _lib.EPI_DGELU) else None
    C2 = torch.empty(M, N, device=DEV, dtype=bf) if epi == _lib.EPI_GELU else None
    f = lambda: ops.gemm(M, N, K, A, A.stride(0), a_k, B, B.stride(0), b_k, Cout, N, epi, C2=C2, aux=aux, accumulate=accumulate)
    for label, (bn, sp, dbg) in cfgs:
        lib.nsb_gemm_debug_config(bn, sp, dbg)
        t = timeit(f)
        lib.nsb_gemm_debug_config(0, 0, 0)
        print(f"{name:24s} {label:28s} {t*1e3:8.1f} us {2.0*M*N*K/t/1e9:7.1f} TF/s", flush=True)


Mt, Cc = 12288, 768
E = _lib
base = [("default", (0, 0, 0)), ("1grp tma", (0, 0, 3)), ("2grp tma", (0, 0, 4)), ("4grp regstore", (0, 0, 7)), ("single-cta", (0, 0, 5))]
run("fc fwd gelu", Mt, 4 * Cc, Cc, True, True, E.EPI_GELU, cfgs=base)
run("fc fwd store-only", Mt, 4 * Cc, Cc, True, True, E.EPI_STORE, cfgs=base)
run("proj dgrad dgelu", Mt, 4 * Cc, Cc, True, False, E.EPI_DGELU, cfgs=base)
run("attn proj res", Mt, Cc, Cc, True, True, E.EPI_RESIDUAL, cfgs=[base[0], base[4]])
run("mlp proj res", Mt, Cc, 4 * Cc, True, True, E.EPI_RESIDUAL, cfgs=[base[0], base[4]])
run("qkv fwd", Mt, 3 * Cc, Cc, True, True, E.EPI_STORE, cfgs=[base[0], base[2], base[4]])
sp = [("default", (0, 0, 0)), ("single-cta", (0, 0, 5))]
run("qkv wgrad", 3 * Cc, Cc, Mt, False, False, E.EPI_ACCUM, True, cfgs=sp)
run("attn proj wgrad", Cc, Cc, Mt, False, False, E.EPI_ACCUM, True, cfgs=sp)
run("fc wgrad", 4 * Cc, Cc, Mt, False, False, E.EPI_ACCUM, True, cfgs=sp)
run("mlp proj wgrad", Cc, 4 * Cc, Mt, False, False, E.EPI_ACCUM, True, cfgs=sp)
