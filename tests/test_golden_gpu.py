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
