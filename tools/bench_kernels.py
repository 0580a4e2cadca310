"""Per-kernel timing on the GPU box (CUDA events, L2 flushed between iterations): our kernels next
to the stock torch op on the cfg4 (GPT-2 124M, B=12, T=1024) shapes.  Not a bench.py metric."""
import json
import math
import sys

import torch
import torch.nn.functional as F

from nanosandbox_b200 import _lib, ops

DEV = "cuda"
PEAKS = json.load(open("MEASURED_PEAKS.json")) if __import__("os").path.exists("MEASURED_PEAKS.json") else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        # keep the GPU busy (~100 us spin) while the host enqueues event, kernel, event: the interval then holds
        # the kernel and not the host's launch path (ctypes for ours, the ATen dispatcher for torch)
        torch.cuda._sleep(200_000)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def gemm_row(name, M, N, K, a_k, b_k, epi=_lib.EPI_STORE, accumulate=False):
    bf = torch.bfloat16
    A = torch.randn((M, K) if a_k else (K, M), device=DEV).to(bf)
    B = torch.randn((N, K) if b_k else (K, N), device=DEV).to(bf)
    cdt = torch.float32 if epi in (_lib.EPI_ACCUM, _lib.EPI_RESIDUAL) else bf
    Cout = torch.zeros(M, N, device=DEV, dtype=cdt)
    aux = torch.randn(M, N, device=DEV, dtype=torch.float32 if epi == _lib.EPI_RESIDUAL else bf) if epi in (_lib.EPI_RESIDUAL, _lib.EPI_DGELU) else None
    C2 = torch.empty(M, N, device=DEV, dtype=bf) if epi == _lib.EPI_GELU else None
    f = lambda: ops.gemm(M, N, K, A, A.stride(0), a_k, B, B.stride(0), b_k, Cout, N, epi, C2=C2, aux=aux, accumulate=accumulate)
    t = timeit(f)
    At = A if a_k else A.t()
    Bt = B.t() if b_k else B
    t_ref = timeit(lambda: torch.matmul(At, Bt))
    fl = 2.0 * M * N * K
    print(f"{name:28s} {M:6d}x{N:6d}x{K:6d} ours {t*1e3:8.1f} us {fl/t/1e9:7.1f} TF/s ({fl/t/1e9/PEAKS['bf16_tflops']*100:5.1f}% of measured) | cuBLAS {t_ref*1e3:8.1f} us {fl/t_ref/1e9:7.1f} TF/s", flush=True)


def clocks():
    import subprocess
    try:
        return subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active",
                               "--format=csv,noheader"], capture_output=True, text=True, timeout=20).stdout.strip()
    except Exception as exc:
        return f"nvidia-smi unavailable ({exc})"


def main():
    Mt, Cc, V = 12288, 768, 50304
    print(f"# {torch.cuda.get_device_name(0)}; torch {torch.__version__}; clocks before: {clocks()}")
    print("# every row: median of 10, L2 flushed before each iteration, same process; cuBLAS = torch.matmul on the same operands "
          "(plain product, no epilogue)")
    E = _lib
    gemm_row("qkv fwd", Mt, 3 * Cc, Cc, True, True)
    gemm_row("attn proj fwd (+res)", Mt, Cc, Cc, True, True, E.EPI_RESIDUAL)
    gemm_row("fc fwd (store)", Mt, 4 * Cc, Cc, True, True)
    gemm_row("fc fwd (+gelu)", Mt, 4 * Cc, Cc, True, True, E.EPI_GELU)
    gemm_row("mlp proj fwd (+res)", Mt, Cc, 4 * Cc, True, True, E.EPI_RESIDUAL)
    gemm_row("lm_head fwd", Mt, V, Cc, True, True)
    gemm_row("attn proj dgrad", Mt, Cc, Cc, True, False)
    gemm_row("mlp proj dgrad (store)", Mt, 4 * Cc, Cc, True, False)
    gemm_row("mlp proj dgrad (+dgelu)", Mt, 4 * Cc, Cc, True, False, E.EPI_DGELU)
    gemm_row("fc dgrad", Mt, Cc, 4 * Cc, True, False)
    gemm_row("qkv dgrad", Mt, Cc, 3 * Cc, True, False)
    gemm_row("lm_head dgrad", Mt, Cc, V, True, False)
    gemm_row("qkv wgrad", 3 * Cc, Cc, Mt, False, False, E.EPI_ACCUM, True)
    gemm_row("attn proj wgrad", Cc, Cc, Mt, False, False, E.EPI_ACCUM, True)
    gemm_row("fc wgrad", 4 * Cc, Cc, Mt, False, False, E.EPI_ACCUM, True)
    gemm_row("mlp proj wgrad", Cc, 4 * Cc, Mt, False, False, E.EPI_ACCUM, True)
    gemm_row("lm_head wgrad", V, Cc, Mt, False, False, E.EPI_ACCUM, True)

    print(f"# clocks after the GEMM rows: {clocks()}")
    if "--gemm-only" in sys.argv:
        return
    # attention
    B, T, H, D = 12, 1024, 12, 64
    qkv = torch.randn(B, T, 3 * Cc, device=DEV).to(torch.bfloat16)
    y = torch.empty(B, T, Cc, device=DEV, dtype=torch.bfloat16)
    dy = torch.randn_like(y)
    lse = torch.empty(B, H, T, device=DEV)
    dqkv = torch.empty_like(qkv)
    delta = torch.empty(B, H, T, device=DEV)
    dq_acc = torch.zeros(B * T, Cc, device=DEV)
    t = timeit(lambda: ops.attention_fwd(qkv, y, lse, B, T, H, D))
    fl = 4.0 * B * H * T * T * D / 2
    print(f"attention fwd  {t*1e3:8.1f} us {fl/t/1e9:7.1f} TF/s (causal flops)")
    t = timeit(lambda: ops.attention_bwd(qkv, y, dy, lse, dqkv, delta, B, T, H, D, dq_acc))
    print(f"attention bwd  {t*1e3:8.1f} us {2.5*fl/t/1e9:7.1f} TF/s (2.5x fwd flops)")
    q, k, v = [z.view(B, T, H, D).transpose(1, 2) for z in qkv.split(Cc, 2)]
    t = timeit(lambda: F.scaled_dot_product_attention(q, k, v, is_causal=True))
    print(f"torch sdpa fwd {t*1e3:8.1f} us {fl/t/1e9:7.1f} TF/s")

    # memory-bound kernels
    x = torch.randn(Mt, Cc, device=DEV)
    w = torch.ones(Cc, device=DEV)
    yb = torch.empty(Mt, Cc, device=DEV, dtype=torch.bfloat16)
    mean = torch.empty(Mt, device=DEV); rstd = torch.empty(Mt, device=DEV)
    t = timeit(lambda: ops.layernorm_fwd(x, w, None, yb, mean, rstd))
    by = Mt * Cc * 6
    print(f"layernorm fwd  {t*1e3:8.1f} us {by/t/1e6:7.1f} GB/s ({by/t/1e6/PEAKS['hbm_gbs']*100:5.1f}% of measured)")
    dyb = torch.randn(Mt, Cc, device=DEV).to(torch.bfloat16)
    dres = torch.randn(Mt, Cc, device=DEV); dx = torch.empty_like(dres); dxl = torch.empty_like(dyb)
    dw = torch.zeros(Cc, device=DEV)
    scratch = torch.empty(ops.layernorm_bwd_scratch_bytes(Mt, Cc) // 4, device=DEV)
    t = timeit(lambda: ops.layernorm_bwd(dyb, x, w, mean, rstd, dres, dx, dxl, dw, None, True, scratch))
    by = Mt * Cc * (2 + 4 + 4 + 4 + 2)
    print(f"layernorm bwd  {t*1e3:8.1f} us {by/t/1e6:7.1f} GB/s ({by/t/1e6/PEAKS['hbm_gbs']*100:5.1f}% of measured)")
    logits = torch.randn(Mt, V, device=DEV).to(torch.bfloat16)
    tg = torch.randint(0, V, (Mt,), device=DEV)
    rl = torch.empty(Mt, device=DEV); ls = torch.empty(Mt, device=DEV); st = torch.empty(2, device=DEV)
    t = timeit(lambda: ops.cross_entropy_fwd(logits, tg, ls, rl, st, V))
    by = Mt * V * 2
    print(f"cross-ent fwd  {t*1e3:8.1f} us {by/t/1e6:7.1f} GB/s ({by/t/1e6/PEAKS['hbm_gbs']*100:5.1f}% of measured)")
    t = timeit(lambda: ops.cross_entropy_bwd(logits, tg, ls, st, None, V))
    by = Mt * V * 4
    print(f"cross-ent bwd  {t*1e3:8.1f} us {by/t/1e6:7.1f} GB/s ({by/t/1e6/PEAKS['hbm_gbs']*100:5.1f}% of measured)")
    hh = torch.randn(Mt, 4 * Cc, device=DEV).to(torch.bfloat16); aa = torch.empty_like(hh)
    t = timeit(lambda: ops.gelu_fwd(hh, aa))
    by = Mt * 4 * Cc * 4
    print(f"gelu fwd       {t*1e3:8.1f} us {by/t/1e6:7.1f} GB/s ({by/t/1e6/PEAKS['hbm_gbs']*100:5.1f}% of measured)")
    t = timeit(lambda: ops.gelu_bwd(aa, hh))
    by = Mt * 4 * Cc * 6
    print(f"gelu bwd       {t*1e3:8.1f} us {by/t/1e6:7.1f} GB/s ({by/t/1e6/PEAKS['hbm_gbs']*100:5.1f}% of measured)")
    n = 124_400_640 // 1024 * 1024
    p = torch.randn(n, device=DEV); g = torch.randn(n, device=DEV); m = torch.zeros(n, device=DEV); v = torch.zeros(n, device=DEV)
    lp = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    fl_ = torch.ones(n // 1024, dtype=torch.uint8, device=DEV)
    t = timeit(lambda: ops.adamw_step(p, g, m, v, lp, fl_, 0.9, 0.95, 1e-8, 0.9999, 1e-3, 0.5, None))
    by = n * 30
    print(f"adamw          {t*1e3:8.1f} us {by/t/1e6:7.1f} GB/s ({by/t/1e6/PEAKS['hbm_gbs']*100:5.1f}% of measured)")
    sc = torch.empty(ops.grad_norm_scratch_bytes(n) // 4, device=DEV); out = torch.empty(2, device=DEV)
    t = timeit(lambda: ops.grad_clip_coef(g, 1.0, sc, out))
    by = n * 4
    print(f"grad norm      {t*1e3:8.1f} us {by/t/1e6:7.1f} GB/s ({by/t/1e6/PEAKS['hbm_gbs']*100:5.1f}% of measured)")


if __name__ == "__main__":
    main()
