"""Host-side logic that needs no GPU: arena layout and buckets, configurator, data streams and
get_batch bit-exactness against the oracle, lr schedule, golden fixtures vs the oracle."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from nanosandbox_b200 import data as ndata
from nanosandbox_b200.arena import CHUNK, ParamArena, bucket_ranges, ready_order
from nanosandbox_b200.model import GPT, GPTConfig
from oracle import train_ref
from oracle.model import GPT as RefGPT, GPTConfig as RefConfig

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("bias", [False, True])
def test_arena_layout_and_buckets(bias):
    m = GPT(GPTConfig(n_layer=3, n_head=2, n_embd=64, block_size=16, vocab_size=65, bias=bias), verbose=False)
    named = ready_order(list(m.named_parameters()), 3)
    assert named[0][0] == "transformer.ln_f.weight" and named[-1][0] == "transformer.wte.weight"
    assert named[-2][0] == "transformer.wpe.weight"
    arena = ParamArena(named, "cpu", with_lowp=False)
    assert arena.numel % CHUNK == 0
    total = 0
    for name, p in named:
        s = arena.slots[name]
        assert s.offset % CHUNK == 0 and s.offset >= total
        total = s.offset + s.numel
        assert s.decay == (p.dim() >= 2)
    flags = arena.chunk_decay
    for s in arena.slots.values():
        assert bool(flags[s.offset // CHUNK]) == s.decay
    ranges = bucket_ranges(arena, 3, bias)
    assert len(ranges) == 4 and ranges[0][0] == 0 and ranges[-1][1] == arena.numel
    s_wte = arena.slots["transformer.wte.weight"]
    assert ranges[-1][0] <= s_wte.offset  # tied wte rides in the last bucket
    assert arena.slots["transformer.ln_f.weight"].offset < ranges[0][1]


def test_same_rng_consumption_and_get_batch_bit_exact():
    kw = dict(n_layer=2, n_head=2, n_embd=64, block_size=32, vocab_size=65, bias=False, dropout=0.0)
    stream_a = train_ref.make_uniform_stream(65, 1 << 14, seed=3)
    stream_b = ndata.uniform_stream(65, 1 << 14, seed=3)
    assert np.array_equal(stream_a, stream_b)
    assert np.array_equal(train_ref.make_markov_stream(65, 5000, 2), ndata.markov_stream(65, 5000, 2))
    torch.manual_seed(1337)
    RefGPT(RefConfig(**kw), verbose=False)
    xa, ya = train_ref.get_batch(stream_a, 12, 32)
    torch.manual_seed(1337)
    GPT(GPTConfig(**kw), verbose=False)
    xb, yb = ndata.get_batch(stream_b, 12, 32, "cpu", pin=False)
    assert xa.dtype == torch.int64 and torch.equal(xa, xb) and torch.equal(ya, yb)
    assert torch.equal(xa[:, 1:], ya[:, :-1])


def test_golden_fixtures_match_the_oracle(golden_dir):
    z = np.load(os.path.join(golden_dir, "tiny_step.npz"))
    kw = dict(n_layer=2, n_head=2, n_embd=64, block_size=32, vocab_size=65, bias=True, dropout=0.0)
    m = RefGPT(RefConfig(**kw), verbose=False)
    with torch.no_grad():
        for n, p in m.named_parameters():
            p.copy_(torch.from_numpy(z["param/" + n]))
    logits, loss = m(torch.from_numpy(z["idx"]), torch.from_numpy(z["tgt"]))
    loss.backward()
    assert abs(loss.item() - float(z["loss"])) < 1e-6
    assert np.allclose(logits.detach().numpy(), z["logits"], atol=1e-5)
    for n, p in m.named_parameters():
        assert np.allclose(p.grad.numpy(), z["grad/" + n], atol=1e-6), n
    c = np.load(os.path.join(golden_dir, "curve_cfg1.npz"))
    assert c["losses"].shape == (200,) and c["losses"][-1] < 0.5 * c["losses"][0]
    assert np.array_equal(c["data_head"], ndata.markov_stream(65, 1 << 16, 0)[:4096])


def _run_cfg(tmp_path, argv):
    script = tmp_path / "probe.py"
    script.write_text(
        "batch_size = 12\nlearning_rate = 6e-4\ndataset = 'openwebtext'\ncompile = True\n"
        f"exec(open({os.path.join(ROOT, 'configurator.py')!r}).read())\n"
        "print('RESULT', batch_size, learning_rate, dataset, compile)\n")
    return subprocess.run([sys.executable, str(script), *argv], capture_output=True, text=True)


def test_configurator_contract(tmp_path):
    cfg = tmp_path / "c.py"
    cfg.write_text("batch_size = 64\ndataset = 'shakespeare_char'\n")
    r = _run_cfg(tmp_path, [str(cfg), "--learning_rate=1e-3", "--compile=False"])
    assert r.returncode == 0, r.stderr
    assert "RESULT 64 0.001 shakespeare_char False" in r.stdout
    assert "Overriding: compile = False" in r.stdout
    assert _run_cfg(tmp_path, ["--nonexistent=3"]).returncode != 0        # unknown key -> error
    assert _run_cfg(tmp_path, ["--batch_size=abc"]).returncode != 0       # type must match
    assert _run_cfg(tmp_path, ["--batch_size=1.5"]).returncode != 0
    assert _run_cfg(tmp_path, ["--compile"]).returncode != 0


def test_train_py_refuses_cpu(tmp_path):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "train.py"), "--device=cpu", "--compile=False",
                        f"--out_dir={tmp_path}"], capture_output=True, text=True, cwd=tmp_path,
                       env={**os.environ, "PYTHONPATH": ROOT})
    assert r.returncode != 0 and "CUDA" in (r.stderr + r.stdout)


def test_entrypoint_derives_node_rank(tmp_path):
    fake = tmp_path / "torchrun"
    fake.write_text("#!/bin/bash\necho \"$@\"\necho NODE_RANK=$NODE_RANK MASTER_ADDR=$MASTER_ADDR\n")
    fake.chmod(0o755)
    env = {**os.environ, "PATH": f"{tmp_path}:{os.environ['PATH']}", "HOSTNAME": "train-multipod-5",
           "NNODES": "8", "NPROC_PER_NODE": "1"}
    env.pop("NODE_RANK", None)
    env.pop("MASTER_ADDR", None)
    r = subprocess.run(["bash", os.path.join(ROOT, "deploy", "entrypoint.sh"), "train.py", "cfg.py"],
                       capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr
    assert "--nnodes=8 --node_rank=5 --nproc_per_node=1" in r.stdout
    assert "MASTER_ADDR=train-multipod-0.train-mp-headless" in r.stdout
    env2 = {**env, "NNODES": "1", "NPROC_PER_NODE": "8"}
    r = subprocess.run(["bash", os.path.join(ROOT, "deploy", "entrypoint.sh"), "train.py"],
                       capture_output=True, text=True, env=env2)
    assert "--standalone --nproc_per_node=8 train.py" in r.stdout


def test_from_hf_state_dict_mapping():
    """HF GPT-2 (random init, no download) -> our GPT: same logits as the oracle given the same weights."""
    tr = pytest.importorskip("transformers")
    torch.manual_seed(0)
    hf_cfg = tr.GPT2Config(vocab_size=97, n_positions=32, n_embd=64, n_layer=2, n_head=2,
                           activation_function="gelu", resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)
    hf = tr.GPT2LMHeadModel(hf_cfg).eval()
    args = dict(n_layer=2, n_head=2, n_embd=64, vocab_size=97, block_size=32, bias=True)
    ours = GPT.from_hf_state_dict(hf.state_dict(), args)
    ref = RefGPT(RefConfig(**args), verbose=False).eval()
    ref.load_state_dict(ours.state_dict())
    idx = torch.randint(0, 97, (2, 32))
    with torch.no_grad():
        a, _ = ref(idx, idx)
        b = hf(idx).logits
    assert torch.allclose(a, b, atol=1e-5, rtol=1e-5)


def test_bench_workloads_flops_and_shared_config():
    """bench.py's workload table: FLOPs per token = SURVEY.md 8d's figures, and both arms of a comparison emit the
    same `config` object."""
    import importlib.util
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench_module", os.path.join(root, "bench.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    assert b.flops_per_token(b.WORKLOADS["gpt2"]["cfg"]) == 854_770_176
    assert b.flops_per_token(b.WORKLOADS["char"]["cfg"]) == 70_958_592
    assert b.flops_per_token(b.WORKLOADS["gpt2-medium"]["cfg"]) == 2_423_298_048
    for name, wl in b.WORKLOADS.items():
        for world in (1, 8):
            total = wl["accum_total"] or world
            c = b.bench_config(name, wl, world, total // world, total * wl["batch"] * wl["cfg"]["block_size"], "bf16")
            assert c["parallelism"] == f"dp{world}" and c["micro_steps_per_rank"] * world == total
            assert c["grad_allreduce"] == ("bf16" if world > 1 else "none")
    assert b.WORKLOADS["gpt2"]["accum_total"] * b.WORKLOADS["gpt2"]["batch"] * 1024 == 491_520
