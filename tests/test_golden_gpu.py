"""CUDA path against the committed golden vectors (tests/golden, generated from the oracle by
tools/make_golden.py): one step known-answer test, data sampling bit-exactness, and the 200-step
loss curve of BASELINE.json configs[0] within 1 % (fp32 and bf16)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
if not torch.cuda.is_available():
    pytest.skip("needs a GPU", allow_module_level=True)

from nanosandbox_b200 import data as ndata  # noqa: E402
from nanosandbox_b200.model import GPT, GPTConfig  # noqa: E402

DEV = "cuda"


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30)


def test_tiny_step_known_answer(golden_dir):
    z = np.load(os.path.join(golden_dir, "tiny_step.npz"))
    kw = dict(n_layer=2, n_head=2, n_embd=64, block_size=32, vocab_size=65, bias=True, dropout=0.0)
    m = GPT(GPTConfig(**kw), verbose=False)
    sd = {k[len("param/"):]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param/")}
    sd["lm_head.weight"] = sd["transformer.wte.weight"]
    m.load_state_dict(sd)
    m.to(DEV)
    m.zero_grad()
    logits, loss = m(torch.from_numpy(z["idx"]).to(DEV), torch.from_numpy(z["tgt"]).to(DEV))
    assert rel(logits.cpu().numpy(), z["logits"]) < 1e-5
    assert abs(loss.item() - float(z["loss"])) < 1e-5 * float(z["loss"])
    loss.backward()
    torch.cuda.synchronize()
    for n, p in m.named_parameters():
        assert rel(p.grad.cpu().numpy(), z["grad/" + n]) < 1e-5, n


def get_lr(it, lr=1e-3, warm=10, decay=200, min_lr=1e-4):
    import math
    if it < warm:
        return lr * (it + 1) / (warm + 1)
    if it > decay:
        return min_lr
    r = (it - warm) / (decay - warm)
    return min_lr + 0.5 * (1.0 + math.cos(math.pi * r)) * (lr - min_lr)


@pytest.mark.parametrize("dtype", ["float32", "bfloat16"])
def test_loss_curve_200_steps_within_one_percent(golden_dir, dtype):
    c = np.load(os.path.join(golden_dir, "curve_cfg1.npz"))
    stream = ndata.markov_stream(65, 1 << 16, 0)
    assert np.array_equal(stream[:4096], c["data_head"])
    torch.manual_seed(1337)
    m = GPT(GPTConfig(n_layer=4, n_head=4, n_embd=128, block_size=64, vocab_size=65, bias=False, dropout=0.0),
            verbose=False).to(DEV)
    opt = m.configure_optimizers(0.1, 1e-3, (0.9, 0.99), "cuda", verbose=False)
    X, Y = ndata.get_batch(stream, 12, 64, DEV)
    assert np.array_equal(X.cpu().numpy(), c["first_x"][0])      # data sampling: bit-exact
    ctx = torch.autocast("cuda", dtype=torch.bfloat16, enabled=(dtype == "bfloat16"))
    losses = []
    for it in range(200):
        for g in opt.param_groups:
            g["lr"] = get_lr(it)
        with ctx:
            _, loss = m(X, Y)
        X, Y = ndata.get_batch(stream, 12, 64, DEV)
        if it < 3:
            assert np.array_equal(X.cpu().numpy(), c["first_x"][it + 1])
        loss.backward()
        opt.clip_grad_norm_(1.0)
        opt.step()
        opt.zero_grad(set_to_none=True)
        losses.append(loss.item())
    losses = np.array(losses)
    dev = np.abs(losses - c["losses"]) / c["losses"]
    tol = 0.01
    assert dev.max() < tol, f"{dtype}: max loss deviation {dev.max():.4f} at step {dev.argmax()}"
    if dtype == "float32":
        psum = float(sum(p.detach().double().sum() for p in m.parameters()))
        assert abs(psum - float(c["param_sum"])) < 1e-3 * abs(float(c["param_sum"]))


def test_loss_curve_char_config_200_steps_vs_oracle_same_gpu():
    """BASELINE.json configs[1] (6L/6H/384, block 256, batch 64, bf16 autocast) for 200 optimizer steps on the
    learnable Markov stream: our step against the oracle model driven by stock torch (CUDA ops, torch AdamW,
    clip_grad_norm_) on the same GPU, same seed, same batches.  north_star: per-step losses within 1 %."""
    from oracle.model import GPT as RefGPT, GPTConfig as RefConfig
    kw = dict(n_layer=6, n_head=6, n_embd=384, block_size=256, vocab_size=65, bias=False, dropout=0.0)
    stream = ndata.markov_stream(65, 1 << 18, 0)
    steps, B, T = 200, 64, 256

    def lr_at(it):
        return get_lr(it, lr=1e-3, warm=20, decay=200, min_lr=1e-4)

    def run(ours: bool):
        torch.manual_seed(1337)
        if ours:
            m = GPT(GPTConfig(**kw), verbose=False).to(DEV)
            opt = m.configure_optimizers(0.1, 1e-3, (0.9, 0.99), "cuda", verbose=False)
        else:
            m = RefGPT(RefConfig(**kw), verbose=False).to(DEV)
            opt = m.configure_optimizers(0.1, 1e-3, (0.9, 0.99), "cuda", verbose=False)
        losses = []
        X, Y = ndata.get_batch(stream, B, T, DEV)
        for it in range(steps):
            for g in opt.param_groups:
                g["lr"] = lr_at(it)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                _, loss = m(X, Y)
            X, Y = ndata.get_batch(stream, B, T, DEV)
            loss.backward()
            if ours:
                opt.clip_grad_norm_(1.0)
            else:
                torch.nn.utils.clip_grad_norm_(m.parameters(), 1.0)
            opt.step()
            opt.zero_grad(set_to_none=True)
            losses.append(loss.item())
        return np.array(losses)

    ref = run(False)
    got = run(True)
    assert ref[-1] < ref[0] - 1.0                      # the stream is learnable: the curve actually moves
    dev = np.abs(got - ref) / ref
    assert dev.max() < 0.01, f"max loss deviation {dev.max():.4f} at step {dev.argmax()} ({got[dev.argmax()]:.4f} vs {ref[dev.argmax()]:.4f})"
