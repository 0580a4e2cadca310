import sys, torch, torch.nn.functional as F
from nanosandbox_b200 import _lib, ops
sys.path.insert(0, "tools")
from bench_kernels import timeit, DEV
lib = _lib.load(); bf = torch.bfloat16; E = _lib
Mt, Cc = 12288, 768
def run(name, M, N, K, epi, dbg=0):
    A = torch.randn(M, K, device=DEV).to(bf); B = torch.randn(N, K, device=DEV).to(bf)
    Cout = torch.zeros(M, N, device=DEV, dtype=bf)
    aux = torch.randn(M, N, device=DEV, dtype=bf) if epi == E.EPI_DGELU else None
    C2 = torch.empty(M, N, device=DEV, dtype=bf) if epi == E.EPI_GELU else None
    lib.nsb_gemm_debug_config(0, 0, dbg)
    t = timeit(lambda: ops.gemm(M, N, K, A, K, True, B, K, True, Cout, N, epi, C2=C2, aux=aux))
    lib.nsb_gemm_debug_config(0, 0, 0)
    print(f"{name:34s} K={K:5d} {t*1e3:8.1f} us", flush=True)
for K in (64, 768):
    run("store", Mt, 4 * Cc, K, E.EPI_STORE)
    run("gelu 2grp tma", Mt, 4 * Cc, K, E.EPI_GELU, 4)
    run("gelu 4grp reg", Mt, 4 * Cc, K, E.EPI_GELU, 7)
    run("dgelu 2grp", Mt, 4 * Cc, K, E.EPI_DGELU, 4)
    run("dgelu 4grp reg", Mt, 4 * Cc, K, E.EPI_DGELU, 7)
h = torch.randn(Mt, 4 * Cc, device=DEV).to(bf)
print("torch F.gelu bf16 [12288,3072]: %.1f us" % (timeit(lambda: F.gelu(h)) * 1e3))
print("torch copy   bf16 [12288,3072]: %.1f us" % (timeit(lambda: h.clone()) * 1e3))
