"""GPU check of nsb_layernorm_bwd_noalloc, the LayerNorm backward for hosts that cannot allocate.

Some users of this library only ever hold their operator's input and output buffers (a plug-in
inside another runtime), or are not allowed to call cudaMalloc where they call us (during a stream
capture, under a fixed memory plan).  They cannot supply the scratch buffer nsb_layernorm_bwd asks
for, so this variant keeps its scratch in an array compiled into the library.  The test drives it
through the C-ABI twice in a row on different shapes -- the second call finds the first call's
leftovers in the shared array -- and expects the same results as the caller-scratch entry point,
to the tolerance the LayerNorm tests use.  It also checks that a row wider than the built-in
scratch is refused with a message.  The file name sorts last so a -x run has finished every other
test before this one."""
import pytest
import torch

pytestmark = pytest.mark.gpu
if not torch.cuda.is_available():
    pytest.skip("needs a GPU", allow_module_level=True)

from nanosandbox_b200 import _lib  # noqa: E402

DEV = "cuda"


def _rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-30)).item()


def _run_pair(M, C, lowp, bias, seed):
    lib = _lib.load()
    torch.manual_seed(seed)
    dt = torch.bfloat16 if lowp else torch.float32
    code = _lib.NSB_BF16 if lowp else _lib.NSB_F32
    x = torch.randn(M, C, device=DEV) * 1.5 + 0.25
    w = torch.randn(C, device=DEV)
    mean = x.mean(-1).contiguous()
    rstd = (x.var(-1, unbiased=False) + 1e-5).rsqrt().contiguous()
    dy = torch.randn(M, C, device=DEV).to(dt)
    dres = torch.randn(M, C, device=DEV)
    st = torch.cuda.current_stream().cuda_stream

    def outputs():
        return (torch.empty(M, C, device=DEV), torch.empty(M, C, device=DEV, dtype=torch.bfloat16),
                torch.full((C,), 0.5, device=DEV), torch.full((C,), -1.5, device=DEV) if bias else None)

    dx0, dl0, dw0, db0 = outputs()
    nbytes = lib.nsb_layernorm_bwd_scratch_bytes(M, C)
    scratch = torch.empty(nbytes // 4, device=DEV)
    rc = lib.nsb_layernorm_bwd(dy.data_ptr(), code, x.data_ptr(), w.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                               dres.data_ptr(), dx0.data_ptr(), dl0.data_ptr(), dw0.data_ptr(),
                               db0.data_ptr() if bias else None, 1, scratch.data_ptr(), nbytes, M, C, st)
    _lib.check(rc, "layernorm_bwd")
    dx1, dl1, dw1, db1 = outputs()
    rc = lib.nsb_layernorm_bwd_noalloc(dy.data_ptr(), code, x.data_ptr(), w.data_ptr(), mean.data_ptr(),
                                       rstd.data_ptr(), dres.data_ptr(), dx1.data_ptr(), dl1.data_ptr(),
                                       dw1.data_ptr(), db1.data_ptr() if bias else None, 1, M, C, st)
    _lib.check(rc, "layernorm_bwd_noalloc")
    torch.cuda.synchronize()
    assert torch.isfinite(dx1).all() and torch.isfinite(dw1).all()
    assert _rel(dx1, dx0) < 1e-5
    assert _rel(dl1, dl0) < 4e-3
    assert _rel(dw1, dw0) < 2e-5
    if bias:
        assert _rel(db1, db0) < 2e-5


def test_noalloc_matches_caller_scratch_entry_point_twice_in_a_row():
    # a wide, tall problem first, then a narrow short one that reuses (part of) the same array
    _run_pair(M=4099, C=1024, lowp=True, bias=True, seed=0)
    _run_pair(M=517, C=384, lowp=False, bias=False, seed=1)


def test_noalloc_rejects_rows_wider_than_the_builtin_scratch():
    lib = _lib.load()
    M, C = 8, 2048
    z = torch.zeros(M, C, device=DEV)
    v = torch.zeros(C, device=DEV)
    r = torch.zeros(M, device=DEV)
    rc = lib.nsb_layernorm_bwd_noalloc(z.data_ptr(), _lib.NSB_F32, z.data_ptr(), v.data_ptr(), r.data_ptr(),
                                       r.data_ptr(), None, z.data_ptr(), None, v.data_ptr(), None, 0, M, C,
                                       torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.nsb_last_error().decode()
    assert "1024" in msg and "built-in scratch" in msg
