"""train.py end to end on the GPU, the way the reference's notebook / Jobs call it
(/root/reference/notebooks/colab_nanoGPT_companion.ipynb:70-79, 107-116): config file + --key=value
overrides, synthetic data written on first use, eval + checkpoint, --init_from=resume, and the
torchrun DDP launch (2 ranks when 2 GPUs are present)."""
import os
import re
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
if not torch.cuda.is_available():
    pytest.skip("needs a GPU", allow_module_level=True)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COMMON = ["config/train_shakespeare_char.py", "--eval_interval=20", "--eval_iters=5", "--log_interval=5",
          "--block_size=128", "--batch_size=16", "--n_layer=2", "--n_head=2", "--n_embd=128", "--dropout=0.0",
          "--compile=False", "--synthetic_kind=markov", "--always_save_checkpoint=True", "--warmup_iters=5",
          "--lr_decay_iters=60"]


def run(cmd, cwd, timeout=600):
    env = {**os.environ, "PYTHONPATH": ROOT}
    r = subprocess.run(cmd, cwd=cwd, env=env, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


def prepare(tmp_path):
    for f in ("train.py", "configurator.py", "model.py", "sample.py"):
        os.symlink(os.path.join(ROOT, f), tmp_path / f)
    os.symlink(os.path.join(ROOT, "config"), tmp_path / "config")


def losses(out):
    return [float(m) for m in re.findall(r"iter \d+: loss ([0-9.]+)", out)]


def test_single_gpu_train_eval_checkpoint_resume(tmp_path):
    prepare(tmp_path)
    out = run([sys.executable, "train.py", *COMMON, "--max_iters=40", "--out_dir=out"], tmp_path)
    assert "number of parameters" in out and "num decayed parameter tensors: 10" in out
    assert "step 20: train loss" in out and "saving checkpoint to out" in out
    ls = losses(out)
    assert ls[-1] < ls[0] - 0.3, ls            # the Markov stream is learnable
    assert os.path.exists(tmp_path / "out" / "ckpt.pt")
    ck = torch.load(tmp_path / "out" / "ckpt.pt", map_location="cpu")
    assert "transformer.h.1.attn.c_attn.weight" in ck["model"] and ck["iter_num"] == 40
    assert set(ck["optimizer"]["state"][0].keys()) == {"step", "exp_avg", "exp_avg_sq"}
    out2 = run([sys.executable, "train.py", *COMMON, "--max_iters=50", "--out_dir=out", "--init_from=resume"], tmp_path)
    assert "Resuming training from out" in out2
    ls2 = losses(out2)
    assert ls2 and ls2[0] < ls[0] - 0.3        # picks up where it left off
    # sample.py on that checkpoint (the synthetic stream's meta.pkl has no character table: prompt and samples are ids)
    out_s = run([sys.executable, "sample.py", "--out_dir=out", "--num_samples=2", "--max_new_tokens=24", "--top_k=10",
                 "--start=1 2 3"], tmp_path)
    assert out_s.count("---------------") == 2
    # float32 (exact parity mode) runs through the same script
    out3 = run([sys.executable, "train.py", *COMMON, "--max_iters=10", "--out_dir=out32", "--dtype=float32"], tmp_path)
    assert losses(out3)


def test_char_recipe_with_its_default_dropout(tmp_path):
    """config/train_shakespeare_char.py as shipped (dropout = 0.2, the recipe both reference runs start from):
    trains, evaluates without dropout, and learns the Markov stream."""
    prepare(tmp_path)
    common = [a for a in COMMON if not a.startswith("--dropout")]
    out = run([sys.executable, "train.py", *common, "--max_iters=60", "--out_dir=out_drop"], tmp_path)
    ls = losses(out)
    assert ls[-1] < ls[0] - 0.3, ls
    ev = [float(m) for m in re.findall(r"step \d+: train loss ([0-9.]+)", out)]
    assert ev and ev[-1] < ev[0] - 0.3, ev      # evaluation loss (model.eval(): no dropout) falls as well
    va = [float(m) for m in re.findall(r"val loss ([0-9.]+)", out)]
    assert va and va[-1] < va[0] - 0.3, va      # ... and on the validation stream (another walk of the same chain)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_torchrun_two_ranks(tmp_path):
    prepare(tmp_path)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
               "--master-addr=127.0.0.1", f"--master-port={port}", "train.py", *COMMON, "--max_iters=30",
               "--gradient_accumulation_steps=4", "--out_dir=out_ddp"], tmp_path)
    assert "tokens per iteration will be: 8,192" in out
    ls = losses(out)
    assert ls[-1] < ls[0] - 0.2, ls
