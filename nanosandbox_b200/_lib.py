"""ctypes binding of the C-ABI in include/nanosandbox_b200.h.

The library is the product: there is no Python/torch fallback.  If it is missing or a call
fails, this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libnanosandbox_b200.so")

NSB_F32, NSB_BF16 = 0, 1
EPI_STORE, EPI_GELU, EPI_RESIDUAL, EPI_ACCUM, EPI_DGELU = 0, 1, 2, 3, 4


class NsbError(RuntimeError):
    pass


class GemmArgs(C.Structure):
    _fields_ = [
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("A", C.c_void_p), ("lda", C.c_int), ("a_kmajor", C.c_int),
        ("B", C.c_void_p), ("ldb", C.c_int), ("b_kmajor", C.c_int),
        ("in_dtype", C.c_int), ("epilogue", C.c_int),
        ("C", C.c_void_p), ("ldc", C.c_int), ("c_dtype", C.c_int),
        ("C2", C.c_void_p), ("aux", C.c_void_p), ("bias", C.c_void_p), ("alpha", C.c_void_p),
        ("accumulate", C.c_int),
    ]


_P, _I, _F, _Z, _U64 = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_ulonglong

# name -> (restype, argtypes); every symbol include/nanosandbox_b200.h declares
SIGNATURES = {
    "nsb_version": (_I, []),
    "nsb_last_error": (C.c_char_p, []),
    "nsb_launch_count": (C.c_ulonglong, []),
    "nsb_embedding_fwd": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "nsb_embedding_bwd": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "nsb_layernorm_fwd": (_I, [_P, _P, _P, _P, _I, _P, _P, _I, _I, _F, _P]),
    "nsb_layernorm_bwd_scratch_bytes": (_Z, [_I, _I]),
    "nsb_layernorm_bwd": (_I, [_P, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _P, _Z, _I, _I, _P]),
    "nsb_layernorm_bwd_noalloc": (_I, [_P, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "nsb_cross_entropy_fwd": (_I, [_P, _I, _P, _P, _P, _P, _I, _I, _I, _P]),
    "nsb_cross_entropy_fwd_bwd": (_I, [_P, _I, _P, _P, _P, _P, _I, _I, _I, _P]),
    "nsb_cross_entropy_bwd": (_I, [_P, _I, _P, _P, _P, _P, _I, _I, _I, _P]),
    "nsb_grad_norm_scratch_bytes": (_Z, [_Z]),
    "nsb_grad_clip_coef": (_I, [_P, _Z, _F, _P, _Z, _P, _P]),
    "nsb_adamw_step": (_I, [_P, _P, _P, _P, _P, _P, _Z, _F, _F, _F, _F, _F, _F, _P, _P]),
    "nsb_cast_f32_to_bf16": (_I, [_P, _P, _Z, _F, _P]),
    "nsb_cast_bf16_to_f32": (_I, [_P, _P, _Z, _P]),
    "nsb_scale_f32": (_I, [_P, _Z, _F, _P]),
    "nsb_gelu_fwd_bf16": (_I, [_P, _P, _Z, _P]),
    "nsb_gelu_bwd_bf16": (_I, [_P, _P, _Z, _P]),
    "nsb_colsum": (_I, [_P, _I, _I, _P, _I, _I, _I, _P]),
    "nsb_gemm": (_I, [C.POINTER(GemmArgs), _P]),
    "nsb_attention_fwd": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "nsb_attention_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "nsb_dropout": (_I, [_P, _I, _P, _P, _I, _Z, _F, _U64, _U64, _P]),
    "nsb_attention_fwd_dropout": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _F, _U64, _U64, _P, _Z, _P]),
    "nsb_attention_bwd_dropout": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _F, _U64, _U64, _P, _Z, _P]),
}
# test / tuning hooks exported by the library but not part of the reference-facing surface
DEBUG_SIGNATURES = {
    "nsb_gemm_debug_config": (None, [_I, _I, _I]),
    "nsb_gemm_debug_timing": (None, [_P]),
    "nsb_gemm_debug_pdl": (None, [_I]),
    "nsb_gemm_debug_pipeline": (None, [_I, _I, _I]),
    "nsb_gemm_debug_scheduler": (None, [_I]),
    "nsb_gemm_debug_store_hint": (None, [_I]),
    "nsb_attention_debug_config": (None, [_I]),
    "nsb_attention_debug_timing": (None, [_P]),
}

_lib = None
GEMM_SCHED_MODE = "auto"


def load() -> C.CDLL:
    """dlopen the library and attach prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NsbError(
            f"{LIB_PATH} is missing: build it with `python -m nanosandbox_b200.build` "
            "(there is no CPU / PyTorch fallback for the kernels)")
    lib = C.CDLL(LIB_PATH)
    for table in (SIGNATURES, DEBUG_SIGNATURES):
        for name, (res, args) in table.items():
            fn = getattr(lib, name)  # AttributeError if a declared symbol is not exported
            fn.restype = res
            fn.argtypes = args
    if os.environ.get("NSB_ATTN_MODE"):                # A/B switch: see nsb_attention_debug_config in attention.cu
        lib.nsb_attention_debug_config(int(os.environ["NSB_ATTN_MODE"]))
    if os.environ.get("NSB_GEMM_PDL", "1") == "0":  # A/B switch for the dependent-launch overlap
        lib.nsb_gemm_debug_pdl(0)
    if os.environ.get("NSB_GEMM_STORE_HINT", "0") == "1":   # A/B switch: evict_first hint on stores of outputs larger than L2
        lib.nsb_gemm_debug_store_hint(1)
    # GEMM tile scheduling: "auto" (default) = fixed-stride walk, switched to the per-launch tile counter by the
    # model while a gradient all-reduce shares the GPU with backward (see GPT._engine_backward); "1" / "0" pin
    # the tile counter / the fixed walk for a whole run
    global GEMM_SCHED_MODE
    GEMM_SCHED_MODE = os.environ.get("NSB_GEMM_DYNAMIC", "auto")
    lib.nsb_gemm_debug_scheduler(1 if GEMM_SCHED_MODE == "1" else 0)
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().nsb_last_error().decode(errors="replace")
        raise NsbError(f"{what} failed (rc={rc}): {msg}")
