"""The C-ABI library loads without a GPU and exports every symbol include/nanosandbox_b200.h
declares; argument validation that needs no device works; the product refuses to run on CPU."""
import os
import re

import pytest
import torch

from nanosandbox_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "nanosandbox_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nsb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in the header but not exported"
        assert s in _lib.SIGNATURES, f"{s} has no ctypes prototype"
    assert set(_lib.SIGNATURES) == set(syms)
    assert lib.nsb_version() >= 100


def test_gemm_args_struct_matches_header_layout():
    # field order/types of nsb_gemm_args: ints and pointers only, 8-byte aligned pointers
    names = [f[0] for f in _lib.GemmArgs._fields_]
    assert names == ["M", "N", "K", "A", "lda", "a_kmajor", "B", "ldb", "b_kmajor", "in_dtype", "epilogue",
                     "C", "ldc", "c_dtype", "C2", "aux", "bias", "alpha", "accumulate"]
    import ctypes
    assert ctypes.sizeof(_lib.GemmArgs) == 112


def test_argument_errors_are_reported_without_a_device():
    lib = _lib.load()
    rc = lib.nsb_layernorm_fwd(None, None, None, None, 0, None, None, 4, 7, 1e-5, None)  # C % 4 != 0
    assert rc != 0 and b"n_embd" in lib.nsb_last_error()
    rc = lib.nsb_adamw_step(None, None, None, None, None, None, 1000, 0.9, 0.95, 1e-8, 1.0, 1e-3, 1.0, None, None)
    assert rc != 0 and b"1024" in lib.nsb_last_error()
    rc = lib.nsb_attention_fwd(None, None, None, 1, 1, 8, 2, 48, None)
    assert rc != 0 and b"head dim" in lib.nsb_last_error()


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only behaviour")
def test_product_has_no_cpu_fallback():
    from nanosandbox_b200.model import GPT, GPTConfig
    m = GPT(GPTConfig(n_layer=1, n_head=2, n_embd=64, block_size=16, vocab_size=65), verbose=False)
    idx = torch.zeros(1, 8, dtype=torch.long)
    with pytest.raises(_lib.NsbError):
        m(idx, idx)
    with pytest.raises(_lib.NsbError):
        m.configure_optimizers(0.1, 1e-3, (0.9, 0.95), "cpu")


def test_no_product_module_imports_the_oracle():
    pkg = os.path.join(ROOT, "nanosandbox_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
    for f in ("train.py", "model.py", "configurator.py"):
        assert not re.search(r"^\s*(from|import)\s+oracle", open(os.path.join(ROOT, f)).read(), flags=re.M), f
