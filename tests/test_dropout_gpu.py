"""Dropout > 0 (GPU): the shakespeare_char recipe trains with dropout = 0.2
(/root/reference/notebooks/colab_nanoGPT_companion.ipynb:71).  torch's dropout masks cannot be reproduced
(different RNG streams), so parity is checked two ways:

* statistics of the kernels (keep rate, scaling, determinism, independence of streams), and
* EXACT parity with the masks injected into the oracle: our masks are a pure function of (seed, stream id,
  element index), so the test regenerates every mask of a forward by running nsb_dropout over ones and gives
  them to an oracle model whose four dropout sites multiply by a given mask instead of drawing one.  Loss and
  every gradient then have to agree to the usual tolerances (fp32 1e-5, bf16 2e-2), forward and backward.
"""
import math

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
if not torch.cuda.is_available():
    pytest.skip("needs a GPU", allow_module_level=True)

from nanosandbox_b200 import ops  # noqa: E402
from nanosandbox_b200.model import GPT, GPTConfig  # noqa: E402
from oracle.model import GPT as RefGPT, GPTConfig as RefConfig  # noqa: E402

DEV = "cuda"


def rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


def mask_of(shape, p, seed, sid, dtype=torch.float32):
    """keep / (1 - p) for every element of a tensor of this shape (linear element order)."""
    ones = torch.ones(shape, device=DEV, dtype=dtype)
    out = torch.empty_like(ones)
    ops.dropout(ones, out, p, seed, sid)
    return out


@pytest.mark.parametrize("p", [0.1, 0.2, 0.5])
@pytest.mark.parametrize("n", [1 << 20, 1000003])
def test_dropout_kernel_statistics(p, n):
    x = torch.randn(n, device=DEV)
    y = torch.empty_like(x)
    ops.dropout(x, y, p, 1234, 7)
    kept = y != 0
    rate = kept.float().mean().item()
    assert abs(rate - (1 - p)) < 4 * math.sqrt(p * (1 - p) / n) + 1e-6, rate      # 4 sigma
    assert rel(y[kept], x[kept] * (1.0 / (1.0 - p))) < 1e-6                        # kept values scaled by 1/(1-p)
    y2 = torch.empty_like(x)
    ops.dropout(x, y2, p, 1234, 7)
    assert torch.equal(y, y2)                                                      # same (seed, stream): same mask
    ops.dropout(x, y2, p, 1234, 8)
    other = y2 != 0
    agree = (kept == other).float().mean().item()                                  # independent streams
    assert abs(agree - ((1 - p) ** 2 + p ** 2)) < 0.01
    ops.dropout(x, y2, p, 1235, 7)
    assert abs(((y2 != 0) == kept).float().mean().item() - ((1 - p) ** 2 + p ** 2)) < 0.01
    # mean is preserved (E[y] = x): sum over many elements of a constant
    c = torch.ones(n, device=DEV)
    ops.dropout(c, y, p, 99, 0)
    assert abs(y.mean().item() - 1.0) < 4 * math.sqrt(p / (1 - p) / n) + 1e-6


def test_dropout_dtypes_residual_inplace_and_backward_mask():
    p, seed, sid = 0.2, 42, 3
    n = 4096 * 6 + 2                                        # ragged tail
    x = torch.randn(n, device=DEV)
    m = mask_of((n,), p, seed, sid)
    res = torch.randn(n, device=DEV)
    y = torch.empty(n, device=DEV)
    ops.dropout(x, y, p, seed, sid, residual=res)
    assert rel(y, res + x * m) < 1e-6
    xb = x.to(torch.bfloat16)
    ops.dropout(xb, y, p, seed, sid, residual=res)           # bf16 in: dropped value rounded to bf16, fp32 add
    assert rel(y, res + (xb.float() * m).to(torch.bfloat16).float()) < 1e-6
    yb = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    ops.dropout(x, yb, p, seed, sid)                          # fp32 gradient -> masked bf16 GEMM operand
    assert rel(yb, (x * m).to(torch.bfloat16)) < 1e-6
    z = x.clone()
    ops.dropout(z, z, p, seed, sid)                           # in place
    assert torch.equal(z, x * m)
    # backward of y = dropout(x) is dy * mask with the SAME mask
    xr = x.clone().requires_grad_(True)
    (xr * m).backward(res)
    g = torch.empty(n, device=DEV)
    ops.dropout(res, g, p, seed, sid)
    assert torch.equal(g, xr.grad)


def _attention_with_mask(qkv, B, T, H, D, mask):
    Cc = H * D
    q, k, v = qkv.split(Cc, dim=2)
    q = q.view(B, T, H, D).transpose(1, 2).float()
    k = k.view(B, T, H, D).transpose(1, 2).float()
    v = v.view(B, T, H, D).transpose(1, 2).float()
    s = (q @ k.transpose(-1, -2)) / math.sqrt(D)
    s = s.masked_fill(~torch.tril(torch.ones(T, T, dtype=torch.bool, device=qkv.device)), float("-inf"))
    pr = torch.softmax(s, dim=-1) * mask                      # dropout AFTER the softmax normalisation, as SDPA does
    return (pr @ v).transpose(1, 2).contiguous().view(B, T, Cc)


@pytest.mark.parametrize("dt,B,T,H,D,tol,scratch", [
    (torch.float32, 2, 100, 3, 32, 1e-5, False), (torch.float32, 1, 256, 2, 64, 1e-5, False),
    (torch.bfloat16, 2, 130, 4, 64, 2e-2, False), (torch.float32, 1, 77, 2, 96, 1e-5, False),
    (torch.bfloat16, 2, 130, 4, 64, 2e-2, True), (torch.bfloat16, 3, 256, 6, 64, 2e-2, True), (torch.bfloat16, 1, 1000, 2, 64, 2e-2, True)])
def test_attention_dropout_matches_masked_reference(dt, B, T, H, D, tol, scratch):
    """scratch=False: the exact kernels draw the bits themselves; True: keep bytes into a scratch buffer, then the
    mma.sync flash kernels (bf16, head dim 64).  Same masks either way."""
    torch.manual_seed(5)
    p, seed, sid = 0.2, 77, 11
    Cc = H * D
    qkv = torch.randn(B, T, 3 * Cc, device=DEV).to(dt)
    mask = mask_of((B, H, T, T), p, seed, sid)
    causal = torch.tril(torch.ones(T, T, device=DEV))
    rate = ((mask != 0).float() * causal).sum().item() / (causal.sum().item() * B * H)
    assert abs(rate - (1 - p)) < 0.01
    y = torch.empty(B, T, Cc, device=DEV, dtype=dt)
    lse = torch.empty(B, H, T, device=DEV)
    ms = torch.empty(B * H * T * T, dtype=torch.uint8, device=DEV) if scratch else None
    ops.attention_fwd_dropout(qkv, y, lse, B, T, H, D, p, seed, sid, ms)
    if scratch:
        assert torch.equal(ms.view(B, H, T, T) != 0, mask != 0)          # the scratch holds exactly the keep bits
    qr = qkv.float().clone().requires_grad_(True)
    ref = _attention_with_mask(qr, B, T, H, D, mask)
    assert rel(y, ref) < tol
    dy = torch.randn(B, T, Cc, device=DEV).to(dt)
    ref.backward(dy.float())
    dqkv = torch.empty_like(qkv)
    delta = torch.empty(B, H, T, device=DEV)
    if scratch:
        ms.zero_()                                                        # backward regenerates it
    ops.attention_bwd_dropout(qkv, y, dy, lse, dqkv, delta, B, T, H, D, p, seed, sid, ms)
    assert rel(dqkv, qr.grad) < tol
    # p = 0 through the same entry points is plain attention
    ops.attention_fwd_dropout(qkv, y, lse, B, T, H, D, 0.0, seed, sid, ms)
    assert rel(y, _attention_with_mask(qkv.float(), B, T, H, D, torch.ones_like(mask))) < tol


# ---- whole model: the oracle with our masks injected -------------------------------------------------------
class _MaskMul(nn.Module):
    def __init__(self):
        super().__init__()
        self.mask = None

    def forward(self, x):
        return x * self.mask.to(x.dtype).view_as(x)


def _masked_attention_forward(self, x):
    B, T, Cc = x.size()
    qkv = self.c_attn(x)
    y = _attention_with_mask(qkv, B, T, self.n_head, Cc // self.n_head, self.p_mask).to(qkv.dtype)
    return self.resid_dropout(self.c_proj(y))


def _inject_masks(ref, ours, B, T, p, call_index):
    """Give the oracle the masks our forward number `call_index` uses (stream ids: see GPT._engine_forward)."""
    cfg = ours.config
    L, H, Cc = cfg.n_layer, cfg.n_head, cfg.n_embd
    seed = torch.cuda.initial_seed() & ((1 << 63) - 1)
    sid0 = call_index * (3 * L + 1)
    ref.transformer.drop.mask = mask_of((B, T, Cc), p, seed, sid0)
    for l, blk in enumerate(ref.transformer.h):
        blk.attn.p_mask = mask_of((B, H, T, T), p, seed, sid0 + 1 + 3 * l)
        blk.attn.resid_dropout.mask = mask_of((B, T, Cc), p, seed, sid0 + 2 + 3 * l)
        blk.mlp.dropout.mask = mask_of((B, T, Cc), p, seed, sid0 + 3 + 3 * l)


def _build(p, **kw):
    torch.manual_seed(1337)
    ref = RefGPT(RefConfig(bias=False, dropout=p, **kw), verbose=False)
    torch.manual_seed(1337)
    ours = GPT(GPTConfig(bias=False, dropout=p, **kw), verbose=False)
    ref.transformer.drop = _MaskMul()
    for blk in ref.transformer.h:
        blk.attn.resid_dropout = _MaskMul()
        blk.mlp.dropout = _MaskMul()
        blk.attn.forward = _masked_attention_forward.__get__(blk.attn)
    return ref.to(DEV), ours.to(DEV)


@pytest.mark.parametrize("lowp", [False, True])
def test_model_with_dropout_matches_oracle_given_the_same_masks(lowp):
    kw = dict(n_layer=3, n_head=4, n_embd=256, block_size=128, vocab_size=65)
    p, B, T = 0.2, 4, 128
    ref, ours = _build(p, **kw)
    tol = 2e-2 if lowp else 2e-5
    g = torch.Generator().manual_seed(1)
    for call in range(2):                                   # two forwards: the stream ids move on
        idx = torch.randint(0, 65, (B, T), generator=g).to(DEV)
        tgt = torch.randint(0, 65, (B, T), generator=g).to(DEV)
        _inject_masks(ref, ours, B, T, p, call)
        ref.zero_grad(set_to_none=True)
        ours.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=lowp):
            _, rloss = ref(idx, tgt)
            _, oloss = ours(idx, tgt)
        rloss.backward()
        oloss.backward()
        torch.cuda.synchronize()
        assert abs(oloss.item() - rloss.item()) < tol * abs(rloss.item()), (call, oloss.item(), rloss.item())
        for (n1, p1), (_, p2) in zip(ref.named_parameters(), ours.named_parameters()):
            e = rel(p2.grad, p1.grad)
            assert e < tol, f"call {call} {n1}: grad rel err {e:.2e}"
    # evaluation mode: no dropout, and the same loss as a dropout-free oracle
    ours.eval()
    plain = RefGPT(RefConfig(bias=False, dropout=0.0, **kw), verbose=False).to(DEV)
    plain.load_state_dict({k: v for k, v in ours.state_dict().items()})
    with torch.no_grad():
        _, l0 = ours(idx, tgt)
        _, l1 = plain(idx, tgt)
    assert abs(l0.item() - l1.item()) < 1e-5 * abs(l1.item())


def test_dropout_changes_the_loss_and_is_reproducible():
    kw = dict(n_layer=2, n_head=2, n_embd=128, block_size=64, vocab_size=65)
    torch.manual_seed(1337)
    ours = GPT(GPTConfig(bias=False, dropout=0.2, **kw), verbose=False).to(DEV)
    idx = torch.randint(0, 65, (4, 64)).to(DEV)
    with torch.no_grad():
        a = ours(idx, idx)[1].item()
        b = ours(idx, idx)[1].item()                         # next forward, next masks
        ours._dropout_calls = 0
        c = ours(idx, idx)[1].item()                         # same call index again: same masks
        ours.eval()
        d = ours(idx, idx)[1].item()
    assert a != b and a == c and d != a
