"""World-size-2 gloo tests on CPU: (1) pin of the oracle: stock DDP with require_backward_grad_sync
toggling == mean-allreduce of locally accumulated gradients (SURVEY.md 8c pin 4); (2) our bucket
manager's host logic (order, scaling, both payload dtypes) with injected CPU pack/unpack."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(fn, world=2):
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_entry, args=(fn, r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0, f"rank failed with exit code {p.exitcode}"
    return [q.get(timeout=5) for _ in range(world)]


def _entry(fn, rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        q.put(fn(rank, world))
    finally:
        dist.destroy_process_group()


def _oracle_ddp(rank, world):
    from oracle.model import GPT, GPTConfig
    from torch.nn.parallel import DistributedDataParallel as DDP
    torch.manual_seed(1337)  # same init on both ranks; DDP would broadcast rank 0 anyway
    cfg = GPTConfig(n_layer=2, n_head=2, n_embd=32, block_size=16, vocab_size=50, bias=False, dropout=0.0)
    m = GPT(cfg, verbose=False)
    ref = GPT(cfg, verbose=False)
    ref.load_state_dict(m.state_dict())
    d = DDP(m)
    g = torch.Generator().manual_seed(100 + rank)
    accum = 3
    for micro in range(accum):
        x = torch.randint(0, 50, (4, 16), generator=g)
        y = torch.randint(0, 50, (4, 16), generator=g)
        d.require_backward_grad_sync = micro == accum - 1
        _, loss = d(x, y)
        (loss / accum).backward()
        _, l2 = ref(x, y)
        (l2 / accum).backward()
    worst = 0.0
    for p, r in zip(m.parameters(), ref.parameters()):
        t = r.grad.clone()
        dist.all_reduce(t)
        t /= world
        worst = max(worst, (p.grad - t).abs().max().item())
    return worst


def test_oracle_ddp_equals_mean_allreduce_of_accumulated_grads():
    assert max(_spawn(_oracle_ddp)) < 1e-6


def _bucket_manager(rank, world):
    from nanosandbox_b200.ddp import GradBucketManager
    out = {}
    for mode in ("fp32", "bf16"):
        torch.manual_seed(5 + rank)
        n = 1024 * 6
        grads = torch.randn(n)
        local = grads.clone()
        ranges = [(0, 2048), (2048, 3072), (3072, n)]
        mgr = GradBucketManager(
            grads, ranges, world, None, mode,
            pack=lambda g, st, inv: st.copy_((g * inv).to(torch.bfloat16)),
            unpack=lambda st, g: g.copy_(st.float()),
            scale=lambda g, inv: g.mul_(inv))
        for k in range(3):
            mgr.reduce_bucket(k)
        mgr.finish()
        if mode == "fp32":
            expect = local / world
            dist.all_reduce(expect)
        else:
            expect = (local / world).to(torch.bfloat16)
            dist.all_reduce(expect)
            expect = expect.float()
        out[mode] = (grads - expect).abs().max().item()
        out[mode + "_bytes"] = mgr.bytes_reduced
        # out-of-order / missing buckets are caught
        mgr.reduce_bucket(1)
        try:
            mgr.finish()
            out[mode + "_order_check"] = False
        except AssertionError:
            out[mode + "_order_check"] = True
    return out


def test_bucket_manager_mean_allreduce_both_payloads():
    for res in _spawn(_bucket_manager):
        assert res["fp32"] < 1e-6
        assert res["bf16"] < 1e-2
        assert res["fp32_bytes"] == 1024 * 6 * 4 and res["bf16_bytes"] == 1024 * 6 * 2
        assert res["fp32_order_check"] and res["bf16_order_check"]


def _verify_replicas(rank, world):
    """GradBucketManager.verify_replicas: passes on identical gradients, names the bucket that differs."""
    from nanosandbox_b200.ddp import GradBucketManager
    n = 1024 * 6
    ranges = [(0, 2048), (2048, 4096), (4096, n)]
    grads = torch.arange(n, dtype=torch.float32) * 0.5
    bk = GradBucketManager(grads, ranges, world, None, "fp32", scale=lambda g, inv: g.mul_(inv))
    out = {"same": bk.verify_replicas()}
    if rank == 1:
        grads[3000] += 1e-3            # one element of bucket 1 on one rank
    out["differs"] = not bk.verify_replicas()
    out["record"] = bk.verify_failures[-1] if bk.verify_failures else None
    return out


def test_replica_verifier_names_the_bucket():
    res = _spawn(_verify_replicas)
    for r in res:
        assert r["same"] and r["differs"]
        assert r["record"]["buckets"] == [1] and r["record"]["step"] == 2
    assert sorted(len(r["record"]["this_rank_differs_from_min"]) for r in res) == [0, 1]
