"""Kernel-level parity (GPU): every C-ABI entry point against the stock torch op it replaces,
called through the C-ABI (ctypes) on seeded inputs.  fp32 kernels: 1e-5 relative; bf16: 2e-2."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a GPU", allow_module_level=True)

from nanosandbox_b200 import _lib, ops  # noqa: E402

DEV = "cuda"


def rel_err(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


def assert_close(a, b, rtol, what=""):
    e = rel_err(a, b)
    assert e < rtol and torch.isfinite(a.float()).all(), f"{what}: rel err {e:.3e} >= {rtol}"


# ------------------------------------------------------------------------------------------------
def test_embedding_fwd_bit_exact_and_bwd():
    torch.manual_seed(0)
    B, T, Cc, V = 3, 40, 128, 65
    wte = torch.randn(V, Cc, device=DEV)
    wpe = torch.randn(T, Cc, device=DEV)
    idx = torch.randint(0, V, (B, T), device=DEV)
    out = torch.empty(B, T, Cc, device=DEV)
    ops.embedding_fwd(idx, wte, wpe, out)
    ref = F.embedding(idx, wte) + F.embedding(torch.arange(T, device=DEV), wpe)
    assert torch.equal(out, ref)  # bit-exact (north_star)
    dx = torch.randn(B, T, Cc, device=DEV)
    dwte = torch.randn(V, Cc, device=DEV)
    dwpe = torch.randn(T, Cc, device=DEV)
    rwte = dwte.clone().index_add_(0, idx.view(-1), dx.view(-1, Cc))
    rwpe0 = dx.sum(0)
    rwpe1 = dwpe + dx.sum(0)
    d0 = dwpe.clone()
    ops.embedding_bwd(idx, dx, dwte, d0, False)
    assert_close(dwte, rwte, 1e-5, "dwte")
    assert_close(d0, rwpe0, 1e-5, "dwpe")
    d1 = dwpe.clone()
    ops.embedding_bwd(idx, dx, dwte.clone(), d1, True)
    assert_close(d1, rwpe1, 1e-5, "dwpe acc")


@pytest.mark.parametrize("Cc", [64, 128, 384, 768, 1024])
@pytest.mark.parametrize("bias", [False, True])
@pytest.mark.parametrize("lowp", [False, True])
def test_layernorm_fwd_bwd(Cc, bias, lowp):
    torch.manual_seed(1)
    M = 517
    x = torch.randn(M, Cc, device=DEV) * 2 + 0.5
    w = torch.randn(Cc, device=DEV)
    b = torch.randn(Cc, device=DEV) if bias else None
    ydt = torch.bfloat16 if lowp else torch.float32
    y = torch.empty(M, Cc, device=DEV, dtype=ydt)
    mean = torch.empty(M, device=DEV)
    rstd = torch.empty(M, device=DEV)
    ops.layernorm_fwd(x, w, b, y, mean, rstd)
    xr = x.clone().requires_grad_(True)
    wr = w.clone().requires_grad_(True)
    br = b.clone().requires_grad_(True) if bias else None
    ref = F.layer_norm(xr, (Cc,), wr, br, 1e-5)
    assert_close(y, ref, 4e-3 if lowp else 1e-5, "ln fwd")
    assert_close(mean, x.mean(-1), 1e-5, "mean")
    dy = torch.randn(M, Cc, device=DEV).to(ydt)
    dres = torch.randn(M, Cc, device=DEV)
    ref.backward(dy.float())
    dx = torch.empty(M, Cc, device=DEV)
    dxl = torch.empty(M, Cc, device=DEV, dtype=torch.bfloat16) if lowp else None
    dw = torch.full((Cc,), 3.0, device=DEV)
    db = torch.full((Cc,), -2.0, device=DEV) if bias else None
    scratch = torch.empty(ops.layernorm_bwd_scratch_bytes(M, Cc) // 4, device=DEV)
    ops.layernorm_bwd(dy, x, w, mean, rstd, dres, dx, dxl, dw, db, True, scratch)
    assert_close(dx, xr.grad + dres, 1e-5, "ln dx")
    assert_close(dw, wr.grad + 3.0, 2e-5, "ln dw")
    if bias:
        assert_close(db, br.grad - 2.0, 2e-5, "ln db")
    if lowp:
        assert_close(dxl, dx, 4e-3, "ln dx lowp")
    # no incoming residual, overwrite mode
    ops.layernorm_bwd(dy, x, w, mean, rstd, None, dx, None, dw, db, False, scratch)
    assert_close(dx, xr.grad, 1e-5, "ln dx no res")
    assert_close(dw, wr.grad, 2e-5, "ln dw overwrite")


@pytest.mark.parametrize("V,ld", [(65, 65), (65, 72), (1000, 1000), (50304, 50304), (50257, 50264)])
@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_cross_entropy(V, ld, dt):
    torch.manual_seed(2)
    M = 257
    store = (torch.randn(M, ld, device=DEV) * 3).to(dt)
    logits = store[:, :V]
    targets = torch.randint(0, V, (M,), device=DEV)
    targets[5] = -1
    targets[100] = -1
    lse = torch.empty(M, device=DEV)
    rl = torch.empty(M, device=DEV)
    stats = torch.empty(2, device=DEV)
    ops.cross_entropy_fwd(store, targets, lse, rl, stats, V)
    lr = logits.float().clone().requires_grad_(True)
    ref = F.cross_entropy(lr, targets, ignore_index=-1)
    tol = 1e-5 if dt == torch.float32 else 1e-4
    assert abs(stats[0].item() - ref.item()) < tol * max(1.0, abs(ref.item())), (stats, ref)
    assert stats[1].item() == M - 2
    assert_close(lse, torch.logsumexp(logits.float(), -1), 1e-5, "lse")
    gs = torch.tensor([0.25], device=DEV)
    (ref * 0.25).backward()
    ops.cross_entropy_bwd(store, targets, lse, stats, gs, V)
    assert_close(store[:, :V], lr.grad, 1e-5 if dt == torch.float32 else 6e-3, "dlogits")
    if ld > V:
        assert (store[:, V:] == 0).all()
    assert (store[5] == 0).all() and (store[100] == 0).all()


@pytest.mark.parametrize("V,ld", [(1000, 1000), (50257, 50304), (65, 72), (50304, 50304)])
def test_cross_entropy_fused_fwd_bwd(V, ld):
    """One-pass loss + unscaled gradient against torch, and against the two separate passes."""
    torch.manual_seed(12)
    M = 131
    store = (torch.randn(M, ld, device=DEV) * 3).to(torch.bfloat16)
    keep = store.clone()
    targets = torch.randint(0, V, (M,), device=DEV)
    targets[7] = -1
    targets[M - 1] = V - 1
    lse, rl, stats = torch.empty(M, device=DEV), torch.empty(M, device=DEV), torch.empty(2, device=DEV)
    assert ops.cross_entropy_fused_ok(store)
    ops.cross_entropy_fwd_bwd(store, targets, lse, rl, stats, V)
    lr = keep[:, :V].float().clone().requires_grad_(True)
    ref = F.cross_entropy(lr, targets, ignore_index=-1)
    assert abs(stats[0].item() - ref.item()) < 1e-4 * max(1.0, abs(ref.item())), (stats, ref)
    assert stats[1].item() == M - 1
    assert_close(lse, torch.logsumexp(keep[:, :V].float(), -1), 1e-5, "lse")
    ref.backward()
    assert_close(store[:, :V].float() / (M - 1), lr.grad, 6e-3, "dlogits (unscaled / n_valid)")
    if ld > V:
        assert (store[:, V:] == 0).all()
    assert (store[7] == 0).all()
    # the separate passes with gscale = n_valid produce the same unscaled gradient up to one bf16 rounding
    two = keep.clone()
    lse2, rl2, st2 = torch.empty(M, device=DEV), torch.empty(M, device=DEV), torch.empty(2, device=DEV)
    ops.cross_entropy_fwd(two, targets, lse2, rl2, st2, V)
    ops.cross_entropy_bwd(two, targets, lse2, st2, torch.tensor([float(M - 1)], device=DEV), V)
    assert_close(lse, lse2, 1e-5, "lse vs two-pass")
    assert_close(rl, rl2, 1e-5, "row loss vs two-pass")
    assert_close(store.float(), two.float(), 1e-2, "dlogits vs two-pass")


def test_cross_entropy_fused_rejects_what_it_cannot_do():
    M, V = 4, 100
    lse, rl, stats = torch.empty(M, device=DEV), torch.empty(M, device=DEV), torch.empty(2, device=DEV)
    tg = torch.zeros(M, dtype=torch.int64, device=DEV)
    with pytest.raises(_lib.NsbError):
        ops.cross_entropy_fwd_bwd(torch.zeros(M, V, device=DEV), tg, lse, rl, stats, V)            # fp32
    with pytest.raises(_lib.NsbError):
        ops.cross_entropy_fwd_bwd(torch.zeros(M, 200000, device=DEV, dtype=torch.bfloat16), tg, lse, rl, stats, V)
    assert not ops.cross_entropy_fused_ok(torch.zeros(M, 100, device=DEV, dtype=torch.bfloat16))  # ld % 8 != 0


def test_grad_norm_and_adamw_match_torch():
    torch.manual_seed(3)
    n = 1024 * 37
    n_decay = 1024 * 30
    p = torch.randn(n, device=DEV)
    flags = torch.zeros(n // 1024, dtype=torch.uint8, device=DEV)
    flags[: n_decay // 1024] = 1
    pd = p[:n_decay].clone().requires_grad_(True)
    pn = p[n_decay:].clone().requires_grad_(True)
    opt = torch.optim.AdamW([{"params": [pd], "weight_decay": 0.1}, {"params": [pn], "weight_decay": 0.0}],
                            lr=6e-4, betas=(0.9, 0.95))
    m = torch.zeros(n, device=DEV)
    v = torch.zeros(n, device=DEV)
    lowp = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    scratch = torch.empty(ops.grad_norm_scratch_bytes(n) // 4, device=DEV)
    out = torch.empty(2, device=DEV)
    for step in range(1, 6):
        g = torch.randn(n, device=DEV) * (0.01 if step % 2 else 1.0)
        pd.grad = g[:n_decay].clone()
        pn.grad = g[n_decay:].clone()
        tn = torch.nn.utils.clip_grad_norm_([pd, pn], 1.0)
        opt.step()
        ops.grad_clip_coef(g, 1.0, scratch, out)
        assert abs(out[0].item() - tn.item()) < 1e-5 * tn.item()
        lr, b1, b2 = 6e-4, 0.9, 0.95
        bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step
        ops.adamw_step(p, g, m, v, lowp, flags, b1, b2, 1e-8, 1 - lr * 0.1, lr / bc1, math.sqrt(bc2), out[1:])
        ref = torch.cat([pd.detach(), pn.detach()])
        assert_close(p, ref, 1e-6, f"adamw step {step}")
        assert_close(lowp, ref, 4e-3, "bf16 shadow")


def test_gelu_standalone_bf16():
    torch.manual_seed(9)
    h = (torch.randn(1000, 3072, device=DEV) * 2).to(torch.bfloat16)
    a = torch.empty_like(h)
    ops.gelu_fwd(h, a)
    assert_close(a, F.gelu(h.float()), 4e-3, "gelu fwd")
    hx = h.float().clone().requires_grad_(True)
    da = torch.randn_like(h)
    F.gelu(hx).backward(da.float())
    ops.gelu_bwd(da, h)
    assert_close(da, hx.grad, 4e-3, "gelu bwd")


def test_casts_and_colsum():
    torch.manual_seed(4)
    x = torch.randn(100003, device=DEV)
    y = torch.empty(100003, device=DEV, dtype=torch.bfloat16)
    ops.cast_f32_to_bf16(x, y, 0.5)
    assert torch.equal(y, (x * 0.5).to(torch.bfloat16))
    z = torch.empty(100003, device=DEV)
    ops.cast_bf16_to_f32(y, z)
    assert torch.equal(z, y.float())
    w = x.clone()
    ops.scale_f32(w, 0.125)
    assert torch.equal(w, x * 0.125)
    dy = torch.randn(300, 70, device=DEV)
    db = torch.ones(70, device=DEV)
    ops.colsum(dy, db, True)
    assert_close(db, dy.sum(0) + 1, 1e-5, "colsum")


# ------------------------------------------------------------------------------------------------
def _gemm_case(dt, M, N, K, a_k, b_k, epi, seed=0, bias=False, accumulate=False, alpha=None):
    torch.manual_seed(seed)
    A = (torch.randn(M, K, device=DEV) / math.sqrt(K)).to(dt)
    Bm = torch.randn(N, K, device=DEV).to(dt)
    def mn_store(X):  # [K, rows] storage with the row stride padded to 8 elements (TMA: 16 B)
        rows = X.shape[0]
        st = torch.zeros(X.shape[1], (rows + 7) // 8 * 8, device=DEV, dtype=dt)
        st[:, :rows] = X.t()
        return st
    A_st = A if a_k else mn_store(A)
    B_st = Bm if b_k else mn_store(Bm)
    acc = A.float() @ Bm.float().t()
    al = None
    if alpha is not None:
        al = torch.tensor([alpha], device=DEV)
        acc = acc * alpha
    bv = torch.randn(N, device=DEV) if bias else None
    lowp = dt == torch.bfloat16
    rnd = (lambda t: t.to(torch.bfloat16).float()) if lowp else (lambda t: t)
    cdt = dt
    C2 = aux = None
    if epi == _lib.EPI_STORE:
        ref = rnd(acc + (bv if bias else 0))
        Cout = torch.empty(M, N, device=DEV, dtype=cdt)
    elif epi == _lib.EPI_GELU:
        h = rnd(acc + (bv if bias else 0))
        ref = (h, rnd(F.gelu(h)))
        Cout = torch.empty(M, N, device=DEV, dtype=cdt)
        C2 = torch.empty(M, N, device=DEV, dtype=cdt)
    elif epi == _lib.EPI_RESIDUAL:
        aux = torch.randn(M, N, device=DEV)
        ref = aux + rnd(acc + (bv if bias else 0))
        Cout = torch.empty(M, N, device=DEV)
    elif epi == _lib.EPI_ACCUM:
        Cout = torch.randn(M, N, device=DEV)
        ref = (Cout.clone() if accumulate else 0) + acc
    else:
        aux = (torch.randn(M, N, device=DEV)).to(cdt)
        hx = aux.float().clone().requires_grad_(True)
        F.gelu(hx).backward(rnd(acc))
        ref = rnd(hx.grad)
        Cout = torch.empty(M, N, device=DEV, dtype=cdt)
    ops.gemm(M, N, K, A_st, A_st.stride(0), a_k, B_st, B_st.stride(0), b_k, Cout, Cout.stride(0), epi,
             C2=C2, aux=aux, bias=bv, alpha=al, accumulate=accumulate)
    torch.cuda.synchronize()
    tol = 1e-2 if lowp else 2e-5
    if epi == _lib.EPI_GELU:
        assert_close(Cout, ref[0], tol, "gelu pre")
        assert_close(C2, ref[1], tol, "gelu out")
    else:
        assert_close(Cout, ref, tol, f"gemm epi {epi}")


SHAPES = [(128, 128, 64), (256, 256, 128), (300, 200, 136), (1024, 768, 768), (512, 65, 384), (384, 1152, 1024)]


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("M,N,K", SHAPES)
def test_gemm_kmajor_store(dt, M, N, K):
    _gemm_case(dt, M, N, K, True, True, _lib.EPI_STORE)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("a_k,b_k", [(True, False), (False, True), (False, False)])
@pytest.mark.parametrize("M,N,K", [(256, 256, 128), (384, 640, 1024), (768, 2304, 2048)])
def test_gemm_mn_major(dt, a_k, b_k, M, N, K):
    _gemm_case(dt, M, N, K, a_k, b_k, _lib.EPI_STORE)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("epi", [_lib.EPI_GELU, _lib.EPI_RESIDUAL, _lib.EPI_DGELU])
@pytest.mark.parametrize("bias", [False, True])
def test_gemm_epilogues(dt, epi, bias):
    _gemm_case(dt, 640, 768, 256, True, True, epi, bias=bias and epi != _lib.EPI_DGELU, alpha=0.5 if epi == _lib.EPI_DGELU else None)


@pytest.mark.parametrize("epi", [_lib.EPI_STORE, _lib.EPI_GELU, _lib.EPI_RESIDUAL, _lib.EPI_DGELU, _lib.EPI_ACCUM])
@pytest.mark.parametrize("M,N,K", [(8192, 1024, 128), (4000, 1160, 192), (2048, 2304, 64)])
@pytest.mark.parametrize("dynamic", [0, 1])
def test_gemm_tc_epilogues_many_tiles_per_cta(epi, M, N, K, dynamic):
    """> 148 tiles: every CTA walks several tiles (TMEM double buffering, staging-buffer reuse and
    the aux prefetch running ahead across tile boundaries), under both tile schedulers."""
    lib = _lib.load()
    lib.nsb_gemm_debug_scheduler(dynamic)
    try:
        _gemm_case(torch.bfloat16, M, N, K, True, True, epi, bias=epi in (_lib.EPI_STORE, _lib.EPI_GELU, _lib.EPI_RESIDUAL),
                   accumulate=True, alpha=0.75)
    finally:
        lib.nsb_gemm_debug_scheduler(1 if _lib.GEMM_SCHED_MODE == "1" else 0)


@pytest.mark.parametrize("groups", [3, 4])
@pytest.mark.parametrize("epi", [_lib.EPI_STORE, _lib.EPI_GELU, _lib.EPI_RESIDUAL, _lib.EPI_DGELU, _lib.EPI_ACCUM])
def test_gemm_tc_epilogue_group_count(epi, groups):
    """One (3) and two (4) epilogue warp groups produce the same results, ragged edges included."""
    lib = _lib.load()
    lib.nsb_gemm_debug_config(0, 0, groups)
    try:
        _gemm_case(torch.bfloat16, 3000, 1000, 192, True, True, epi, bias=epi in (_lib.EPI_STORE, _lib.EPI_GELU),
                   accumulate=True, alpha=0.5)
        _gemm_case(torch.bfloat16, 8192, 1024, 128, True, False, epi, accumulate=True)
    finally:
        lib.nsb_gemm_debug_config(0, 0, 0)


@pytest.mark.parametrize("epi", [_lib.EPI_STORE, _lib.EPI_GELU, _lib.EPI_RESIDUAL, _lib.EPI_DGELU, _lib.EPI_ACCUM])
def test_gemm_tc_direct_epilogue_fallback(epi):
    """The direct-store epilogue (used when an output row stride is not 16-byte addressable)."""
    lib = _lib.load()
    lib.nsb_gemm_debug_config(0, 0, 2)
    try:
        _gemm_case(torch.bfloat16, 1000, 520, 192, True, True, epi, bias=epi != _lib.EPI_DGELU and epi != _lib.EPI_ACCUM,
                   accumulate=True)
    finally:
        lib.nsb_gemm_debug_config(0, 0, 0)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("accumulate", [False, True])
@pytest.mark.parametrize("M,N,K", [(768, 768, 4096), (2304, 768, 2048), (65, 384, 1024), (200, 120, 520)])
def test_gemm_wgrad_accum(dt, accumulate, M, N, K):
    # weight-gradient form: both operands MN-major, fp32 accumulate output (split-K when accumulating)
    _gemm_case(dt, M, N, K, False, False, _lib.EPI_ACCUM, accumulate=accumulate, alpha=0.125)


# The three lm_head GEMMs of the headline configuration (cfg4: M = 12 x 1024 tokens, C = 768, V = 50304): 31 % of the
# step's GEMM FLOPs.  The two gradient GEMMs read the 1.2 GB dlogits, which is what switches the tile walk
# to n-fastest and picks the split-K plan at M = 50304; no smaller shape reaches those branches by policy.
@pytest.mark.parametrize("which", ["fwd", "dgrad", "wgrad"])
@pytest.mark.parametrize("dynamic", [0, 1])
def test_gemm_lm_head_headline_shapes(which, dynamic):
    T, Cc, V = 12288, 768, 50304
    _lib.load().nsb_gemm_debug_scheduler(dynamic)
    try:
        _lm_head_case(which, T, Cc, V)
    finally:
        _lib.load().nsb_gemm_debug_scheduler(1 if _lib.GEMM_SCHED_MODE == "1" else 0)


def _lm_head_case(which, T, Cc, V):
    if which == "fwd":       # logits[T, V] = lnf[T, C] . wte[V, C]^T
        _gemm_case(torch.bfloat16, T, V, Cc, True, True, _lib.EPI_STORE)
    elif which == "dgrad":   # dlnf[T, C] = alpha * dlogits[T, V] . wte[V, C]   (B read MN-major in place)
        _gemm_case(torch.bfloat16, T, Cc, V, True, False, _lib.EPI_STORE, alpha=1.0 / T)
    else:                    # dwte[V, C] += alpha * dlogits^T[V, T] . lnf[T, C] (both MN-major, split-K, n-fastest walk)
        _gemm_case(torch.bfloat16, V, Cc, T, False, False, _lib.EPI_ACCUM, accumulate=True, alpha=1.0 / T)


@pytest.mark.parametrize("walk", [8, 9])
@pytest.mark.parametrize("epi,a_k,b_k,M,N,K", [
    (_lib.EPI_STORE, True, True, 2048, 1280, 256),
    (_lib.EPI_STORE, True, False, 1536, 768, 1024),
    (_lib.EPI_ACCUM, False, False, 2304, 768, 2048),
    (_lib.EPI_RESIDUAL, True, True, 4000, 1160, 192),
])
def test_gemm_tc_tile_walk_orders(walk, epi, a_k, b_k, M, N, K):
    """m-fastest (8) and n-fastest (9) tile walks forced on shapes with several tiles per CTA."""
    lib = _lib.load()
    lib.nsb_gemm_debug_config(0, 0, walk)
    try:
        _gemm_case(torch.bfloat16, M, N, K, a_k, b_k, epi, accumulate=True, alpha=0.5)
    finally:
        lib.nsb_gemm_debug_config(0, 0, 0)


@pytest.mark.parametrize("epi,a_k,b_k,M,N,K", [
    (_lib.EPI_STORE, True, True, 8192, 1024, 128),
    (_lib.EPI_RESIDUAL, True, True, 4000, 1160, 192),
    (_lib.EPI_DGELU, True, False, 3000, 1000, 192),
    (_lib.EPI_ACCUM, False, False, 2304, 768, 2048),
])
def test_gemm_tc_tile_schedulers(epi, a_k, b_k, M, N, K):
    """Both tile schedulers (0: fixed-stride walk, 1: per-launch tile counter) give the same results."""
    lib = _lib.load()
    try:
        for dynamic in (0, 1):
            lib.nsb_gemm_debug_scheduler(dynamic)
            _gemm_case(torch.bfloat16, M, N, K, a_k, b_k, epi, accumulate=True, alpha=0.5)
    finally:
        lib.nsb_gemm_debug_scheduler(1 if _lib.GEMM_SCHED_MODE == "1" else 0)


def test_gemm_tc_tile_counters_are_recycled():
    """More launches than there are tile counters (1024), on two streams at once: every launch must find its
    counter back at zero, i.e. still compute every tile exactly once."""
    torch.manual_seed(3)
    _lib.load().nsb_gemm_debug_scheduler(1)
    M, N, K = 1024, 768, 256
    A = (torch.randn(M, K, device=DEV) / 16).to(torch.bfloat16)
    Bm = torch.randn(N, K, device=DEV).to(torch.bfloat16)
    ref = (A.float() @ Bm.float().t())
    side = torch.cuda.Stream()
    outs = [torch.zeros(M, N, device=DEV) for _ in range(2)]
    acc = [torch.zeros(M, N, device=DEV) for _ in range(2)]
    torch.cuda.synchronize()
    n = 1300
    for i in range(n):
        ops.gemm(M, N, K, A, K, True, Bm, K, True, outs[0], N)
        ops.gemm(M, N, K, A, K, True, Bm, K, True, acc[0], N, _lib.EPI_ACCUM, accumulate=True)
        with torch.cuda.stream(side):
            ops.gemm(M, N, K, A, K, True, Bm, K, True, outs[1], N)
            ops.gemm(M, N, K, A, K, True, Bm, K, True, acc[1], N, _lib.EPI_ACCUM, accumulate=True)
    torch.cuda.synchronize()
    _lib.load().nsb_gemm_debug_scheduler(1 if _lib.GEMM_SCHED_MODE == "1" else 0)
    for j in range(2):
        assert_close(outs[j], ref, 1e-2, "last store")
        assert_close(acc[j], ref * n, 1e-2, "sum over all launches")   # a lost or repeated tile shows up here


def test_gemm_tc_matches_simt_path():
    """The tcgen05 kernel against our own exact SIMT kernel on identical bf16 inputs."""
    torch.manual_seed(7)
    M, N, K = 512, 512, 512
    A = torch.randn(M, K, device=DEV).to(torch.bfloat16)
    Bm = torch.randn(N, K, device=DEV).to(torch.bfloat16)
    c1 = torch.empty(M, N, device=DEV)
    c2 = torch.empty(M, N, device=DEV)
    lib = _lib.load()
    ops.gemm(M, N, K, A, K, True, Bm, K, True, c1, N)
    lib.nsb_gemm_debug_config(0, 0, 1)
    try:
        ops.gemm(M, N, K, A, K, True, Bm, K, True, c2, N)
    finally:
        lib.nsb_gemm_debug_config(0, 0, 0)
    torch.cuda.synchronize()
    assert_close(c1, c2, 1e-5, "tc vs simt")


# ------------------------------------------------------------------------------------------------
def _attention_ref(qkv, B, T, H, D):
    Cc = H * D
    q, k, v = qkv.split(Cc, dim=2)
    q = q.view(B, T, H, D).transpose(1, 2)
    k = k.view(B, T, H, D).transpose(1, 2)
    v = v.view(B, T, H, D).transpose(1, 2)
    y = F.scaled_dot_product_attention(q, k, v, is_causal=True)
    return y.transpose(1, 2).contiguous().view(B, T, Cc)


@pytest.mark.parametrize("path", [0, 2])
@pytest.mark.parametrize("B,T,H", [(2, 128, 3), (1, 1024, 2), (3, 100, 2), (2, 384, 4), (2, 64, 2), (3, 1000, 12)])
def test_attention_bf16_paths(path, B, T, H):
    """tcgen05 forward (path 0) and the mma.sync flash kernels (path 2) against torch SDPA.  The last shape has
    288 (key block, head) items with a ragged last block: more than one per SM, so the persistent backward walks
    several items per CTA (K / V double buffer, dK / dV hand-over between items, clipped TMA stores)."""
    lib = _lib.load()
    lib.nsb_attention_debug_config(path)
    try:
        test_attention_fwd_bwd(torch.bfloat16, B, T, H, 64)
    finally:
        lib.nsb_attention_debug_config(0)


@pytest.mark.parametrize("mode", [12, 14, 5, 6, 10, 16])
@pytest.mark.parametrize("B,T,H", [(2, 256, 3), (1, 1000, 2)])
def test_attention_tc_kernel_variants(mode, B, T, H):
    """The earlier forms of the tcgen05 kernels stay selectable for A/B runs (8 probability warps backward: 12,
    4 softmax warps forward: 14, every A operand from shared memory: 5, K/V from tensor memory: 6, forward P
    through shared memory: 10; the 4-warp forward is the one modes 10 / 11 act on; 16-warp backward with one CTA
    per (head, key block) instead of the persistent walk: 16): each must pass the same parity check as the defaults."""
    lib = _lib.load()
    if mode in (5, 6):
        lib.nsb_attention_debug_config(12)      # these two switches act on the 8-warp backward kernel
    if mode == 10:
        lib.nsb_attention_debug_config(14)
    lib.nsb_attention_debug_config(mode)
    try:
        test_attention_fwd_bwd(torch.bfloat16, B, T, H, 64)
    finally:
        for m in (13, 15, 9, 11, 17):           # back to the defaults
            lib.nsb_attention_debug_config(m)


@pytest.mark.parametrize("dt,B,T,H,D", [
    (torch.float32, 2, 64, 4, 32), (torch.float32, 2, 100, 2, 64),
    (torch.bfloat16, 2, 256, 6, 64), (torch.bfloat16, 1, 1024, 2, 64), (torch.bfloat16, 3, 100, 2, 64),
    (torch.bfloat16, 2, 64, 4, 32)])
def test_attention_fwd_bwd(dt, B, T, H, D):
    torch.manual_seed(5)
    Cc = H * D
    qkv = (torch.randn(B, T, 3 * Cc, device=DEV)).to(dt)
    dy = torch.randn(B, T, Cc, device=DEV).to(dt)
    y = torch.empty(B, T, Cc, device=DEV, dtype=dt)
    lse = torch.empty(B, H, T, device=DEV)
    ops.attention_fwd(qkv, y, lse, B, T, H, D)
    ref_in = qkv.float().clone().requires_grad_(True)
    ref = _attention_ref(ref_in, B, T, H, D)
    tol = 1e-5 if dt == torch.float32 else 1e-2
    assert_close(y, ref, tol, "attn fwd")
    ref.backward(dy.float())
    dqkv = torch.zeros_like(qkv)
    delta = torch.empty(B, H, T, device=DEV)
    dq_acc = torch.zeros(B * T, Cc, device=DEV) if (dt == torch.bfloat16 and D == 64) else None
    ops.attention_bwd(qkv, y, dy, lse, dqkv, delta, B, T, H, D, dq_acc)
    torch.cuda.synchronize()
    if dq_acc is not None:
        assert not dq_acc.any()  # the accumulator is handed back zeroed
    g = ref_in.grad
    btol = 2e-5 if dt == torch.float32 else 2e-2
    assert_close(dqkv[..., :Cc], g[..., :Cc], btol, "dq")
    assert_close(dqkv[..., Cc:2 * Cc], g[..., Cc:2 * Cc], btol, "dk")
    assert_close(dqkv[..., 2 * Cc:], g[..., 2 * Cc:], btol, "dv")
