"""Known-answer pins for the oracle: the reference ships no golden vectors, so
the restatement is pinned against nanoGPT's own published log lines, ln V, an independently
written GPT-2 (Hugging Face) and stock DDP semantics."""
import math
import os

import pytest
import torch

from oracle.model import GPT, GPTConfig, flops_per_token
from oracle.train_ref import get_lr

CFGS = {
    1: dict(n_layer=4, n_head=4, n_embd=128, block_size=64, vocab_size=65),
    2: dict(n_layer=6, n_head=6, n_embd=384, block_size=256, vocab_size=65),
    4: dict(n_layer=12, n_head=12, n_embd=768, block_size=1024, vocab_size=50304),
    5: dict(n_layer=24, n_head=16, n_embd=1024, block_size=1024, vocab_size=50304),
}
PARAMS = {1: 804_096, 2: 10_745_088, 4: 124_373_760, 5: 354_599_936}
FLOPS = {1: 5_168_640, 2: 70_958_592, 4: 854_770_176, 5: 2_423_298_048}
DECAY = {1: (18, 802_944, 9, 1_152), 2: (26, 10_740_096, 13, 4_992),
         4: (50, 124_354_560, 25, 19_200), 5: (98, 354_549_760, 49, 50_176)}


@pytest.mark.parametrize("k", [1, 2, 4, 5])
def test_param_counts_and_flops(k):
    cfg = GPTConfig(bias=False, dropout=0.0, **CFGS[k])
    with torch.device("meta"):
        m = GPT(cfg, verbose=False)
    assert sum(p.numel() for p in m.parameters()) == PARAMS[k]
    assert flops_per_token(cfg) == FLOPS[k]
    assert 6 * m.get_num_params() + 12 * cfg.n_layer * cfg.n_embd * cfg.block_size == FLOPS[k]
    ps = list(m.parameters())
    dec = [p for p in ps if p.dim() >= 2]
    nod = [p for p in ps if p.dim() < 2]
    assert (len(dec), sum(p.numel() for p in dec), len(nod), sum(p.numel() for p in nod)) == DECAY[k]


def test_nanogpt_log_lines():
    with torch.device("meta"):
        m2 = GPT(GPTConfig(bias=False, **CFGS[2]), verbose=False)
        m4 = GPT(GPTConfig(bias=False, **CFGS[4]), verbose=False)
    assert "%.2fM" % (m2.get_num_params() / 1e6) == "10.65M"
    assert "%.2fM" % (m4.get_num_params() / 1e6) == "123.59M"
    assert len(list(m4.named_parameters())) == 75  # tied lm_head deduplicated


def test_loss_at_init_is_ln_v():
    torch.manual_seed(1337)
    cfg = GPTConfig(bias=False, **CFGS[1])
    m = GPT(cfg, verbose=False)
    x = torch.randint(0, 65, (12, 64))
    y = torch.randint(0, 65, (12, 64))
    _, loss = m(x, y)
    assert abs(loss.item() - math.log(65)) < 0.1


def test_state_dict_keys():
    with torch.device("meta"):
        m = GPT(GPTConfig(bias=True, **CFGS[1]), verbose=False)
    keys = set(m.state_dict().keys())
    assert "transformer.wte.weight" in keys and "lm_head.weight" in keys
    assert "transformer.h.3.attn.c_attn.weight" in keys
    assert "transformer.h.0.mlp.c_proj.bias" in keys
    assert "transformer.ln_f.bias" in keys


def test_logits_equal_hf_gpt2():
    """Independent second oracle: HF GPT2LMHeadModel with exact-erf GELU and the four Conv1D
    weights transposed reproduces the restatement's logits bit for bit on CPU."""
    tr = pytest.importorskip("transformers")
    torch.manual_seed(0)
    cfg = GPTConfig(bias=True, n_layer=2, n_head=2, n_embd=64, block_size=32, vocab_size=97)
    m = GPT(cfg, verbose=False).eval()
    hf_cfg = tr.GPT2Config(vocab_size=97, n_positions=32, n_embd=64, n_layer=2, n_head=2,
                           activation_function="gelu", resid_pdrop=0.0, embd_pdrop=0.0,
                           attn_pdrop=0.0, layer_norm_epsilon=1e-5)
    hf = tr.GPT2LMHeadModel(hf_cfg).eval()
    sd = m.state_dict()
    hsd = hf.state_dict()
    transposed = ("attn.c_attn.weight", "attn.c_proj.weight", "mlp.c_fc.weight", "mlp.c_proj.weight")
    new = {}
    for k in hsd:
        if k.endswith(".attn.bias") or k.endswith(".attn.masked_bias"):
            new[k] = hsd[k]
            continue
        assert k in sd, k
        new[k] = sd[k].t().contiguous() if k.endswith(transposed) else sd[k].clone()
    hf.load_state_dict(new)
    idx = torch.randint(0, 97, (3, 32))
    with torch.no_grad():
        ours, _ = m(idx, idx)
        theirs = hf(idx).logits
    assert torch.allclose(ours, theirs, atol=1e-5, rtol=1e-5), (ours - theirs).abs().max()


def test_get_lr_pinned_form():
    assert get_lr(0, 1e-3, 100, 5000, 1e-4) == pytest.approx(1e-3 * 1 / 101)
    assert get_lr(99, 1e-3, 100, 5000, 1e-4) == pytest.approx(1e-3 * 100 / 101)
    assert get_lr(100, 1e-3, 100, 5000, 1e-4) == pytest.approx(1e-3)
    assert get_lr(5000, 1e-3, 100, 5000, 1e-4) == pytest.approx(1e-4)
    assert get_lr(6000, 1e-3, 100, 5000, 1e-4) == 1e-4
