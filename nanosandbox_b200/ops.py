"""Tensor-level wrappers over the C-ABI.  torch supplies device memory and streams only; every
arithmetic op below is one of our CUDA kernels (include/nanosandbox_b200.h)."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import (EPI_ACCUM, EPI_DGELU, EPI_GELU, EPI_RESIDUAL, EPI_STORE, NSB_BF16, NSB_F32,
                   GemmArgs, check)

_DT = {torch.float32: NSB_F32, torch.bfloat16: NSB_BF16}


def _require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise _lib.NsbError(f"{name} must be a CUDA tensor: nanosandbox_b200 has no CPU path")


def stream_handle() -> int:
    return torch.cuda.current_stream().cuda_stream


def dtype_code(t: torch.Tensor) -> int:
    return _DT[t.dtype]


def ptr(t):
    return None if t is None else t.data_ptr()


def embedding_fwd(idx, wte, wpe, out, stream=None):
    _require_cuda(idx, "idx")
    B, T = idx.shape
    V, Cc = wte.shape
    check(_lib.load().nsb_embedding_fwd(idx.data_ptr(), wte.data_ptr(), wpe.data_ptr(), out.data_ptr(),
                                        B, T, Cc, V, stream or stream_handle()), "embedding_fwd")
    return out


def embedding_bwd(idx, dx, dwte, dwpe, accumulate_wpe, stream=None):
    B, T = idx.shape
    V, Cc = dwte.shape
    check(_lib.load().nsb_embedding_bwd(idx.data_ptr(), dx.data_ptr(), dwte.data_ptr(), dwpe.data_ptr(),
                                        B, T, Cc, V, int(accumulate_wpe), stream or stream_handle()),
          "embedding_bwd")


def layernorm_fwd(x, w, b, y, mean, rstd, eps=1e-5, stream=None):
    _require_cuda(x, "x")
    Cc = x.shape[-1]
    M = x.numel() // Cc
    check(_lib.load().nsb_layernorm_fwd(x.data_ptr(), w.data_ptr(), ptr(b), y.data_ptr(), dtype_code(y),
                                        mean.data_ptr(), rstd.data_ptr(), M, Cc, eps,
                                        stream or stream_handle()), "layernorm_fwd")
    return y


def layernorm_bwd_scratch_bytes(M, Cc) -> int:
    return int(_lib.load().nsb_layernorm_bwd_scratch_bytes(M, Cc))


def layernorm_bwd(dy, x, w, mean, rstd, dres_in, dx, dx_lowp, dw, db, accumulate, scratch, stream=None):
    Cc = x.shape[-1]
    M = x.numel() // Cc
    check(_lib.load().nsb_layernorm_bwd(dy.data_ptr(), dtype_code(dy), x.data_ptr(), w.data_ptr(),
                                        mean.data_ptr(), rstd.data_ptr(), ptr(dres_in), dx.data_ptr(),
                                        ptr(dx_lowp), dw.data_ptr(), ptr(db), int(accumulate),
                                        scratch.data_ptr(), scratch.numel() * scratch.element_size(),
                                        M, Cc, stream or stream_handle()), "layernorm_bwd")


def cross_entropy_fwd(logits, targets, lse, row_loss, stats, V, stream=None):
    """logits: [M, ld] storage (ld = logits.stride(0)), first V columns valid."""
    _require_cuda(logits, "logits")
    M, ld = logits.shape[0], logits.stride(0)
    check(_lib.load().nsb_cross_entropy_fwd(logits.data_ptr(), dtype_code(logits), targets.data_ptr(),
                                            lse.data_ptr(), row_loss.data_ptr(), stats.data_ptr(), M, V, ld,
                                            stream or stream_handle()), "cross_entropy_fwd")


def cross_entropy_fwd_bwd(logits, targets, lse, row_loss, stats, V, stream=None):
    """Loss statistics and, in place, the UNSCALED gradient softmax - onehot (bf16 logits only)."""
    _require_cuda(logits, "logits")
    M, ld = logits.shape[0], logits.stride(0)
    check(_lib.load().nsb_cross_entropy_fwd_bwd(logits.data_ptr(), dtype_code(logits), targets.data_ptr(),
                                                lse.data_ptr(), row_loss.data_ptr(), stats.data_ptr(), M, V, ld,
                                                stream or stream_handle()), "cross_entropy_fwd_bwd")


def cross_entropy_fused_ok(logits) -> bool:
    """True if nsb_cross_entropy_fwd_bwd accepts this logits buffer (bf16, aligned rows that fit in smem)."""
    ld = logits.stride(0)
    return logits.dtype == torch.bfloat16 and ld % 8 == 0 and ld * 2 <= 232448 - 1024 and logits.data_ptr() % 16 == 0


def cross_entropy_bwd(logits, targets, lse, stats, gscale, V, stream=None):
    M, ld = logits.shape[0], logits.stride(0)
    check(_lib.load().nsb_cross_entropy_bwd(logits.data_ptr(), dtype_code(logits), targets.data_ptr(),
                                            lse.data_ptr(), stats.data_ptr(), ptr(gscale), M, V, ld,
                                            stream or stream_handle()), "cross_entropy_bwd")


def grad_norm_scratch_bytes(n) -> int:
    return int(_lib.load().nsb_grad_norm_scratch_bytes(n))


def grad_clip_coef(g, max_norm, scratch, out, stream=None):
    check(_lib.load().nsb_grad_clip_coef(g.data_ptr(), g.numel(), float(max_norm), scratch.data_ptr(),
                                         scratch.numel() * scratch.element_size(), out.data_ptr(),
                                         stream or stream_handle()), "grad_clip_coef")


def adamw_step(p, g, m, v, p_lowp, chunk_decay, beta1, beta2, eps, decay_factor, step_size, bc2_sqrt,
               gscale, stream=None):
    check(_lib.load().nsb_adamw_step(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), ptr(p_lowp),
                                     chunk_decay.data_ptr(), p.numel(), beta1, beta2, eps, decay_factor,
                                     step_size, bc2_sqrt, ptr(gscale), stream or stream_handle()),
          "adamw_step")


def cast_f32_to_bf16(src, dst, scale=1.0, stream=None):
    check(_lib.load().nsb_cast_f32_to_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), float(scale),
                                           stream or stream_handle()), "cast_f32_to_bf16")


def cast_bf16_to_f32(src, dst, stream=None):
    check(_lib.load().nsb_cast_bf16_to_f32(src.data_ptr(), dst.data_ptr(), src.numel(),
                                           stream or stream_handle()), "cast_bf16_to_f32")


def scale_f32(x, scale, stream=None):
    check(_lib.load().nsb_scale_f32(x.data_ptr(), x.numel(), float(scale), stream or stream_handle()),
          "scale_f32")


def gelu_fwd(h, a, stream=None):
    check(_lib.load().nsb_gelu_fwd_bf16(h.data_ptr(), a.data_ptr(), h.numel(), stream or stream_handle()), "gelu_fwd")


def gelu_bwd(da, h, stream=None):
    """in place: da <- da * gelu'(h)"""
    check(_lib.load().nsb_gelu_bwd_bf16(da.data_ptr(), h.data_ptr(), h.numel(), stream or stream_handle()), "gelu_bwd")


def colsum(dy, db, accumulate, stream=None):
    M, N = dy.shape
    check(_lib.load().nsb_colsum(dy.data_ptr(), dtype_code(dy), dy.stride(0), db.data_ptr(), int(accumulate),
                                 M, N, stream or stream_handle()), "colsum")


def gemm(M, N, K, A, lda, a_kmajor, Bm, ldb, b_kmajor, Cout, ldc, epilogue=EPI_STORE, C2=None, aux=None,
         bias=None, alpha=None, accumulate=False, stream=None):
    """C[M,N] = alpha * sum_k A[m,k] B[n,k] (+ epilogue).  A/B: torch tensors (f32 or bf16)."""
    _require_cuda(A, "A")
    args = GemmArgs(M, N, K, A.data_ptr(), lda, int(a_kmajor), Bm.data_ptr(), ldb, int(b_kmajor),
                    dtype_code(A), epilogue, Cout.data_ptr(), ldc, dtype_code(Cout),
                    ptr(C2), ptr(aux), ptr(bias), ptr(alpha), int(accumulate))
    check(_lib.load().nsb_gemm(C.byref(args), stream or stream_handle()), "gemm")


def attention_fwd(qkv, y, lse, B, T, H, D, stream=None):
    _require_cuda(qkv, "qkv")
    check(_lib.load().nsb_attention_fwd(qkv.data_ptr(), y.data_ptr(), lse.data_ptr(), dtype_code(qkv),
                                        B, T, H, D, stream or stream_handle()), "attention_fwd")


def attention_bwd(qkv, y, dy, lse, dqkv, delta, B, T, H, D, dq_acc=None, stream=None):
    """dq_acc: zero-filled fp32 [B*T, H*D] accumulator (tensor-core path) or None."""
    check(_lib.load().nsb_attention_bwd(qkv.data_ptr(), y.data_ptr(), dy.data_ptr(), lse.data_ptr(),
                                        dqkv.data_ptr(), delta.data_ptr(), ptr(dq_acc), dtype_code(qkv), B, T, H, D,
                                        stream or stream_handle()), "attention_bwd")


def dropout(x, y, p, seed, stream_id, residual=None, stream=None):
    """y = (residual or 0) + dropout(x), mask = f(seed, stream_id, element index); y may be x (in place)."""
    _require_cuda(x, "x")
    check(_lib.load().nsb_dropout(x.data_ptr(), dtype_code(x), ptr(residual), y.data_ptr(), dtype_code(y), x.numel(),
                                  float(p), int(seed), int(stream_id), stream or stream_handle()), "dropout")
    return y


def attention_fwd_dropout(qkv, y, lse, B, T, H, D, p, seed, stream_id, mask_scratch=None, stream=None):
    """mask_scratch: uint8 tensor of >= B*H*T*T elements (bf16, D = 64: the mma.sync kernels) or None (exact kernels)."""
    _require_cuda(qkv, "qkv")
    check(_lib.load().nsb_attention_fwd_dropout(qkv.data_ptr(), y.data_ptr(), lse.data_ptr(), dtype_code(qkv), B, T, H, D,
                                                float(p), int(seed), int(stream_id), ptr(mask_scratch),
                                                0 if mask_scratch is None else mask_scratch.numel(),
                                                stream or stream_handle()), "attention_fwd_dropout")


def attention_bwd_dropout(qkv, y, dy, lse, dqkv, delta, B, T, H, D, p, seed, stream_id, mask_scratch=None, stream=None):
    check(_lib.load().nsb_attention_bwd_dropout(qkv.data_ptr(), y.data_ptr(), dy.data_ptr(), lse.data_ptr(), dqkv.data_ptr(),
                                                delta.data_ptr(), dtype_code(qkv), B, T, H, D, float(p), int(seed),
                                                int(stream_id), ptr(mask_scratch),
                                                0 if mask_scratch is None else mask_scratch.numel(),
                                                stream or stream_handle()), "attention_bwd_dropout")


def gemm_dynamic_tiles(on: bool) -> None:
    """In "auto" mode (NSB_GEMM_DYNAMIC unset): GEMM launches from now on draw their tiles from a per-launch
    counter (on) or walk a fixed stride (off).  The counter form keeps every running CTA busy when some SMs are
    held by another kernel (an NCCL all-reduce beside backward); the fixed walk is ~0.5 % faster when the GEMM
    has the GPU to itself.  No effect when the environment pins one form."""
    if _lib.GEMM_SCHED_MODE == "auto":
        _lib.load().nsb_gemm_debug_scheduler(1 if on else 0)


def launch_count() -> int:
    return int(_lib.load().nsb_launch_count())
