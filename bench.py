"""Headline benchmark: GPT-2 124M data-parallel training throughput (tokens/s, MFU) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # our arm (CUDA kernels)
    python bench.py --impl reference [--gpus N] --steps K --warmup W  # the CPU restatement, host cores

A "step" is one optimizer step of BASELINE.json's headline configuration (cfg4): batch 12 x block
1024 x 40 accumulation micro-steps = 491,520 tokens, split evenly over the ranks (40/N micro-steps
each), synthetic uniform tokens over the 50304 vocabulary, random-init weights, bf16 autocast
policy, grad clip 1.0, AdamW.  Under torchrun (N>1) one rank per GPU, our bucketed NCCL all-reduce.
Prints ONE JSON line on rank 0 (keys: see the bench contract in DESIGN.md section 6).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# BASELINE.json configs -> bench workloads.  "gpt2" (configs[3]) is the one the metric is quoted on and the default;
# "char" is configs[1] at N=1 and configs[2] at N=8 (one 64x256 micro-step per rank, gradient_accumulation_steps =
# world size, as the reference launches it); "gpt2-medium" is configs[4]'s model (its 8-Pod launch form is
# tools/multipod_emulation.sh; under plain torchrun this measures the same step single-Pod).
WORKLOADS = {
    "gpt2": dict(cfg=dict(n_layer=12, n_head=12, n_embd=768, block_size=1024, vocab_size=50304, bias=False, dropout=0.0),
                 batch=12, accum_total=40, scaling="strong", cpu_micro_batch=2,
                 label="gpt2-124M (12L/12H/768, block 1024, vocab 50304) batch 12 x accum 40 = 491520 tokens/step, bf16 [BASELINE.json configs[3]]"),
    "char": dict(cfg=dict(n_layer=6, n_head=6, n_embd=384, block_size=256, vocab_size=65, bias=False, dropout=0.0),
                 batch=64, accum_total=None, scaling="weak", cpu_micro_batch=64,
                 label="shakespeare_char GPT (6L/6H/384, block 256, vocab 65) batch 64, one micro-step per rank "
                       "(gradient_accumulation_steps = world size), bf16 [BASELINE.json configs[1] at N=1, configs[2] at N=8]"),
    "gpt2-medium": dict(cfg=dict(n_layer=24, n_head=16, n_embd=1024, block_size=1024, vocab_size=50304, bias=False, dropout=0.0),
                        batch=12, accum_total=40, scaling="strong", cpu_micro_batch=1,
                        label="gpt2-medium 350M (24L/16H/1024, block 1024, vocab 50304) batch 12 x accum 40 = 491520 tokens/step, bf16 [BASELINE.json configs[4] model]"),
}
NOMINAL_BF16 = 2.25e15


def flops_per_token(cfg) -> int:
    """nanoGPT's estimate_mfu numerator per token: 6N + 12*L*H*Q*T, N = parameters without wpe."""
    L, H, C, T, V = cfg["n_layer"], cfg["n_head"], cfg["n_embd"], cfg["block_size"], cfg["vocab_size"]
    per_block = 12 * C * C + 2 * C + (13 * C if cfg["bias"] else 0)
    n = V * C + L * per_block + C + (C if cfg["bias"] else 0)
    return 6 * n + 12 * L * H * (C // H) * T


def gemm_traffic_bytes(workload: str):
    """(dram bytes per GEMM launch, where the figure comes from).  ncu cannot run inside a timed bench, so the
    figure is the one of the committed `ncu --set full` capture of this same workload under profiles/ (newest
    first); None for workloads without a capture."""
    if workload != "gpt2":
        return None, None
    for name in ("r02_gemm_traffic.json", "r01d_gemm_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                return float(json.load(f)["dram_bytes_per_launch"]), f"profiles/{name} (ncu capture of this workload, not measured in this run)"
        except (OSError, KeyError, ValueError):
            continue
    return None, None


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf=d["bf16_tflops"], tf_sus=d.get("bf16_tflops_sustained", d["bf16_tflops"]), src="measured")
    return dict(hbm=6650.0, tf=1590.0, tf_sus=1400.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm = sorted(int(float(r[0])) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [int(float(r[1])) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def cpu_reference_run(wl, world: int, n_micro: int, warmup: int):
    """The reference's CPU path (oracle restatement, stock torch ops) on all host cores, on a bounded sample of
    the workload.  A full optimizer step is `k` forward+backward micro-steps followed by ONE clip + AdamW, so
    the two parts are timed separately on the sample (`n_micro` micro-steps of `cpu_micro_batch` sequences, two
    optimizer updates) and composed: step time = (tokens per step / tokens per micro-step) * t_micro + t_opt."""
    from oracle.model import GPT as RefGPT, GPTConfig as RefConfig
    cfg = wl["cfg"]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(1337)
    model = RefGPT(RefConfig(**cfg), verbose=False)
    opt = model.configure_optimizers(0.1, 6e-4, (0.9, 0.95), "cpu", verbose=False)
    g = torch.Generator().manual_seed(0)
    toks = torch.randint(0, cfg["vocab_size"], (1 << 20,), generator=g)
    T, mb = cfg["block_size"], wl["cpu_micro_batch"]
    accum_total = wl["accum_total"] or world
    tokens_per_step = accum_total * wl["batch"] * T

    def micro():
        ix = torch.randint(len(toks) - T - 1, (mb,))
        x = torch.stack([toks[i:i + T] for i in ix])
        y = torch.stack([toks[i + 1:i + 1 + T] for i in ix])
        _, loss = model(x, y)
        loss.backward()

    def update():
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        opt.zero_grad(set_to_none=True)

    for _ in range(max(1, warmup)):
        micro()
    update()                                   # warm (allocates the AdamW state)
    t0 = time.perf_counter()
    for _ in range(n_micro):
        micro()
    t_micro = (time.perf_counter() - t0) / n_micro
    t0 = time.perf_counter()
    update()
    t_opt = time.perf_counter() - t0
    k = tokens_per_step / (mb * T)
    step_s = k * t_micro + t_opt
    sample = (f"{n_micro} forward+backward micro-steps of {mb}x{T} tokens ({t_micro * 1e3:.0f} ms each) and one clip+AdamW "
              f"update ({t_opt * 1e3:.0f} ms), fp32, oracle/ restatement on stock torch CPU ops, {cores} threads; composed "
              f"as {k:.0f} micro-steps + 1 update per {tokens_per_step}-token optimizer step")
    return tokens_per_step / step_s, step_s * 1e3, cores, sample, t_micro * 1e3


def torch_gpu_baseline(dev, wl, micro_steps: int = 8):
    """Informational: the SAME step on stock PyTorch CUDA ops (oracle model under bf16 autocast,
    SDPA flash attention, cuBLAS, fused AdamW) on this GPU, for `micro_steps` micro-steps."""
    from oracle.model import GPT as RefGPT, GPTConfig as RefConfig
    torch.manual_seed(1337)
    CFG, BATCH = wl["cfg"], wl["batch"]
    micro_steps = min(micro_steps, wl["accum_total"] or 1)
    model = RefGPT(RefConfig(**CFG), verbose=False).to(dev)
    opt = model.configure_optimizers(0.1, 6e-4, (0.9, 0.95), "cuda", verbose=False)
    T, V = CFG["block_size"], CFG["vocab_size"]
    x = torch.randint(0, V, (BATCH, T), device=dev)
    y = torch.randint(0, V, (BATCH, T), device=dev)

    def step():
        for _ in range(micro_steps):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                _, loss = model(x, y)
            (loss / micro_steps).backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        opt.zero_grad(set_to_none=True)

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(2):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 2
    del model, opt
    torch.cuda.empty_cache()
    return {"value": micro_steps * BATCH * T / (ms / 1e3), "unit": "tokens/s",
            "what": f"stock PyTorch {torch.__version__} eager (cuBLAS, SDPA, fused AdamW, bf16 autocast), same model, "
                    f"{micro_steps} micro-steps per optimizer step, same GPU; informational, not the reference arm"}


def bench_config(name, wl, world, accum, tokens_per_step, bucket_dtype):
    """The `config` object of the result line, identical for both arms of a comparison."""
    return {"workload": wl["label"], "parallelism": f"dp{world}", "micro_steps_per_rank": accum,
            "tokens_per_step": tokens_per_step,
            "l2": ("inputs larger than L2: every micro-step streams several GB of activations/weights (126 MB L2)"
                   if name != "char" else
                   "inputs larger than L2: one micro-step streams ~1 GB of activations (126 MB L2)"),
            "grad_allreduce": bucket_dtype if world > 1 else "none"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = WORKLOADS[args.config]
    world = max(1, args.gpus)
    n_micro = max(1, min(args.steps, 40))          # one sample is about a second of CPU work: K of them, bounded
    warm = max(0, min(args.warmup, 5))
    accum_total = wl["accum_total"] or world
    T = wl["cfg"]["block_size"]
    tok_s, composed_ms, cores, sample, micro_ms = cpu_reference_run(wl, world, n_micro, warm)
    # a "step" of this arm is one timed sample (a forward+backward micro-step): ms_per_step is what was measured;
    # `value` is the throughput of the full optimizer step composed from the samples (see cpu_baseline.sample)
    line = {"impl": "reference", "metric": "train_tokens_per_sec", "value": tok_s, "unit": "tokens/s",
            "n_gpus": args.gpus, "steps": n_micro, "warmup": warm, "ms_per_step": micro_ms,
            "composed_optimizer_step_ms": composed_ms, "higher_is_better": True,
            "scaling": wl["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": bench_config(args.config, wl, world, accum_total // world, accum_total * wl["batch"] * T,
                                   args.bucket_dtype),
            "cpu_baseline": {"value": tok_s, "unit": "tokens/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": tok_s, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


_RESULT_OUT = None


def keep_stdout_clean():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL's version banner, warnings
    from native code), so file descriptor 1 is pointed at stderr for the run and the result line goes to a
    private duplicate of the original stdout."""
    global _RESULT_OUT
    if _RESULT_OUT is None:
        sys.stdout.flush()
        _RESULT_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _RESULT_OUT if _RESULT_OUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="gpt2", choices=sorted(WORKLOADS),
                    help="BASELINE.json workload; the default (gpt2 = configs[3]) is the headline line")
    ap.add_argument("--bucket-dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-torch-baseline", action="store_true")
    args = ap.parse_args()
    keep_stdout_clean()
    if args.impl == "reference":
        return run_reference(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback "
                         "(use --impl reference for the CPU restatement)")
    args.warmup = max(args.warmup, 3)

    import torch.distributed as dist
    from nanosandbox_b200 import ops
    from nanosandbox_b200.data import uniform_stream
    from nanosandbox_b200.ddp import DistributedDataParallel as DDP
    from nanosandbox_b200.model import GPT, GPTConfig

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    ddp = world > 1
    if ddp:
        dist.init_process_group(backend="nccl", device_id=dev)
    wl = WORKLOADS[args.config]
    CFG, BATCH = wl["cfg"], wl["batch"]
    accum_total = wl["accum_total"] or world           # "char": one micro-step per rank
    assert accum_total % world == 0, f"{accum_total} micro-steps do not split over {world} ranks"
    accum = accum_total // world
    T, V = CFG["block_size"], CFG["vocab_size"]
    tokens_per_step = accum_total * BATCH * T
    FLOPS_PER_TOKEN = flops_per_token(CFG)

    torch.manual_seed(1337 + rank)
    model = GPT(GPTConfig(**CFG), verbose=False).to(dev)
    opt = model.configure_optimizers(0.1, 6e-4, (0.9, 0.95), "cuda", verbose=False)
    raw = model
    if ddp:
        model = DDP(model, device_ids=[local_rank], bucket_dtype=args.bucket_dtype)
    data = uniform_stream(V, 1 << 22, seed=0)
    ctx = torch.autocast("cuda", dtype=torch.bfloat16)

    def host_batch():
        ix = torch.randint(len(data) - T, (BATCH,))
        x = torch.stack([torch.from_numpy(data[i:i + T].astype(np.int64)) for i in ix])
        y = torch.stack([torch.from_numpy(data[i + 1:i + 1 + T].astype(np.int64)) for i in ix])
        return x.pin_memory(), y.pin_memory()

    def next_host_batch():
        xh, yh = host_batch()
        return xh.to(dev, non_blocking=True), yh.to(dev, non_blocking=True)

    pending = [None]            # the batch fetched ahead, as nanoGPT's loop does ("prefetch the next batch")

    def train_step(batches, from_host: bool, sync: bool = True, update: bool = True):
        """One optimizer step through the public API (GPT.forward / backward / clip / AdamW).  from_host: the
        loop of train.py -- every micro-step's batch is built on the host, copied from pinned memory, and the
        NEXT one is fetched right after the forward has been enqueued, while the GPU works."""
        loss = None
        if from_host and pending[0] is None:
            pending[0] = next_host_batch()
        for micro in range(accum):
            if ddp:
                model.require_backward_grad_sync = sync and (micro == accum - 1)
            X, Y = pending[0] if from_host else batches[micro]
            with ctx:
                _, loss = model(X, Y)
                loss = loss / accum
            if from_host:
                pending[0] = next_host_batch()
            loss.backward()
        if update:
            opt.clip_grad_norm_(1.0)
            opt.step()
            opt.zero_grad(set_to_none=True)
        return loss

    resident = [tuple(t.to(dev) for t in host_batch()) for _ in range(accum)]

    def barrier():
        if ddp:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if ddp:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item()

    # ---- warm-up (the first step doubles as the sanity check: a random-init model must sit at ln V), then the
    #      device-resident timed region ("value") ----
    loss0 = train_step(resident, False).item() * accum
    loss_tol = 0.35 if CFG["n_embd"] > 768 else 0.2      # wider models start with more logit variance
    if not abs(loss0 - math.log(V)) < loss_tol:
        raise SystemExit(f"bench.py: loss at init {loss0:.4f} is not within {loss_tol} of ln({V}) = {math.log(V):.4f}: "
                         "the step is not computing the model")
    for _ in range(args.warmup - 1):
        train_step(resident, False)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = ops.launch_count()
    if ddp:
        model.buckets.time_comm = False
    ms = timed(lambda: train_step(resident, False), args.steps)
    launches = ops.launch_count() - launches0
    value = tokens_per_step * args.steps / (ms / 1e3)

    # ---- end-to-end: host batches from pinned memory + loss read back every step ----
    def e2e_step():
        loss = train_step(None, True)
        return loss.item()  # device -> host read of the step's result
    e2e_step()
    ms_e2e = timed(e2e_step, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    e2e_value = tokens_per_step * args.steps / (ms_e2e / 1e3)
    h2d = accum * world * 2 * BATCH * T * 8
    d2h = 4 * world

    # ---- host side of one step: how long the Python / ctypes path needs to enqueue it, GPU idle at the start
    #      (for a step with more launches than the driver's queue holds this includes waiting for the GPU) ----
    enq = []
    for _ in range(3):
        barrier()
        t0 = time.perf_counter()
        train_step(resident, False)
        enq.append((time.perf_counter() - t0) * 1e3)
    torch.cuda.synchronize()
    host = {"launches_per_step": launches / args.steps, "enqueue_ms_per_step": min(enq),
            "device_ms_per_step": ms / args.steps}

    # ---- roofline of the dominant kernel (tcgen05 GEMM): CUDA events around every GEMM launch
    #      of one more step, on the launching stream ----
    peaks = load_peaks()
    ev = []
    orig_gemm = ops.gemm

    ext_streams = {}

    def timed_gemm(M, N, K, *a, **k):
        # events go on the stream the GEMM is launched on (the weight-gradient GEMMs run on a side stream)
        h = k.get("stream")
        if h is None or h == torch.cuda.current_stream().cuda_stream:
            strm = torch.cuda.current_stream()
        else:
            strm = ext_streams.setdefault(h, torch.cuda.ExternalStream(h))
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(strm)
        orig_gemm(M, N, K, *a, **k)
        e.record(strm)
        ev.append((2.0 * M * N * K, s, e))
    import nanosandbox_b200.model as model_mod
    model_mod.ops.gemm = timed_gemm
    keep_overlap = raw.overlap_wgrad
    raw.overlap_wgrad = False   # one stream for this pass: each GEMM's events then bracket that GEMM alone
    try:
        train_step(resident, False)
        torch.cuda.synchronize()
    finally:
        model_mod.ops.gemm = orig_gemm
        raw.overlap_wgrad = keep_overlap
    g_flops = sum(f for f, _, _ in ev)
    g_ms = sum(s.elapsed_time(e) for _, s, e in ev)
    achieved = g_flops / (g_ms / 1e3) / 1e12
    roofline = {"kernel": "gemm_tc_kernel (tcgen05/TMEM/TMA bf16 GEMM, all layouts/epilogues)", "bound": "tensor",
                "achieved": achieved, "peak": peaks["tf_sus"], "unit": "TFLOP/s", "frac": achieved / peaks["tf_sus"],
                "traffic": gemm_traffic_bytes(args.config)[0], "traffic_source": gemm_traffic_bytes(args.config)[1],
                "peak_source": f"{peaks['src']} (MEASURED_PEAKS.json bf16_tflops_sustained: kernel timed inside a long step)",
                "launches_timed": len(ev), "avg_launch_us": g_ms * 1e3 / max(1, len(ev)),
                "algorithmic_flops_per_launch": g_flops / max(1, len(ev)),
                "gemm_share_of_step": g_ms / (ms / args.steps)}

    # ---- data-parallel path (N > 1): correctness check, exposed communication, bus bandwidth ----
    comm = None
    ddp_check = None
    if ddp:
        bk = model.buckets
        # (1) replicas: every rank's parameter arena must be bit-identical to rank 0's after the timed steps
        ar = raw.arena()
        ref_params = ar.params.clone()
        dist.broadcast(ref_params, src=0)
        same = torch.tensor([int(torch.equal(ref_params, ar.params))], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        where = None
        if not bool(same.item()):                    # say where the replicas differ: it names the gradient that is not synchronised
            ne = (ref_params != ar.params) | torch.isnan(ar.params)
            idx = ne.nonzero().flatten()
            where = {"rank": rank, "differing": int(idx.numel()), "nan": int(torch.isnan(ar.params).sum().item())}
            if idx.numel():
                slots = {}
                for name, sl in ar.slots.items():
                    n = int(((idx >= sl.offset) & (idx < sl.offset + sl.numel)).sum().item())
                    if n:
                        slots[name] = n
                top = max(slots, key=slots.get) if slots else None
                if top is not None:                  # shape of the region: a run of rows (a collective's chunk) or a tile
                    sl = ar.slots[top]
                    loc = idx[(idx >= sl.offset) & (idx < sl.offset + sl.numel)] - sl.offset
                    ncol = sl.shape[-1]
                    r, c = loc // ncol, loc % ncol
                    where["region"] = {"slot": top, "shape": list(sl.shape), "rows": [int(r.min()), int(r.max())],
                                       "cols": [int(c.min()), int(c.max())], "first": int(loc.min()), "last": int(loc.max())}
                where["max_abs_diff"] = float((ref_params - ar.params)[idx].abs().max().item())
                where["slots"] = dict(sorted(slots.items(), key=lambda kv: -kv[1])[:8])
                where["outside_slots"] = int(idx.numel()) - sum(slots.values())
            print(f"bench.py: rank {rank} parameters differ from rank 0: {where}", file=sys.stderr, flush=True)
        del ref_params
        # (2) one step's synchronised gradients against all_reduce(locally accumulated gradients) / world
        opt.zero_grad(set_to_none=True)
        train_step(resident, False, sync=False, update=False)
        expect = ar.grads.clone()
        dist.all_reduce(expect, op=dist.ReduceOp.SUM)
        expect /= world
        opt.zero_grad(set_to_none=True)
        train_step(resident, False, sync=True, update=False)
        torch.cuda.synchronize()
        gerr = ((ar.grads - expect).norm() / (expect.norm() + 1e-30)).reshape(1)
        dist.all_reduce(gerr, op=dist.ReduceOp.MAX)
        # (3) the synchronised gradients themselves must be the same bits on every rank, step after step
        grads_same = True
        for rep in range(int(os.environ.get("NSB_BENCH_GRAD_REPEATS", "4"))):
            if rep:
                opt.zero_grad(set_to_none=True)
                train_step(resident, False, sync=True, update=False)
                torch.cuda.synchronize()
            ref_g = ar.grads.clone()
            dist.broadcast(ref_g, src=0)
            ne = (ref_g != ar.grads)
            bad = torch.tensor([int(ne.any().item())], device=dev)
            dist.all_reduce(bad, op=dist.ReduceOp.MAX)
            if ne.any().item():
                idx = ne.nonzero().flatten()
                desc = {"rank": rank, "repeat": rep, "differing": int(idx.numel())}
                for name, sl in ar.slots.items():
                    loc = idx[(idx >= sl.offset) & (idx < sl.offset + sl.numel)] - sl.offset
                    if loc.numel():
                        ncol = sl.shape[-1]
                        r, c = loc // ncol, loc % ncol
                        desc.setdefault("slots", {})[name] = {"n": int(loc.numel()), "shape": list(sl.shape), "rows": [int(r.min()), int(r.max())],
                                                              "cols": [int(c.min()), int(c.max())]}
                        if len(desc["slots"]) >= 4:
                            break
                desc["max_abs_diff"] = float((ref_g - ar.grads)[idx].abs().max().item())
                print(f"bench.py: rank {rank} synchronised gradients differ from rank 0: {desc}", file=sys.stderr, flush=True)
            del ref_g
            if bad.item():
                grads_same = False
        gtol = 1e-2 if args.bucket_dtype == "bf16" else 1e-5      # bf16 payload: 8 mantissa bits per addend
        del expect
        opt.zero_grad(set_to_none=True)
        ddp_check = {"params_bit_equal_across_ranks": bool(same.item()), "synced_grads_bit_equal_across_ranks": grads_same,
                     "synced_grad_rel_err": gerr.item(),
                     "tolerance": gtol, "loss_at_init": loss0, "ln_vocab": math.log(V),
                     "ok": bool(same.item()) and grads_same and gerr.item() < gtol}
        if not bool(same.item()):
            # which rank drifted and where ("open issue"): reported in the result line, the
            # throughput numbers of the run stand
            wheres = [None] * world
            dist.all_gather_object(wheres, where)
            ddp_check["replicas_differ"] = [w for w in wheres if w and w.get("differing")]
        if gerr.item() >= gtol or not grads_same:
            # the synchronised gradients are wrong or differ between ranks: the data-parallel path is broken
            raise SystemExit(f"bench.py: data-parallel check failed: {ddp_check}")
        if not ddp_check["ok"]:
            print(f"bench.py: replicas are not bit-identical after the timed steps: {ddp_check}", file=sys.stderr, flush=True)

        # (3) time on the communication stream (pack + all-reduce + unpack, waiting included) and the
        #     stand-alone bus bandwidth of the same payload
        bk.time_comm = True
        bk.comm_events = []
        b0 = bk.bytes_reduced
        train_step(resident, False)
        torch.cuda.synchronize()
        t_ms = sum(a.elapsed_time(b) for a, b in bk.comm_events)
        nbytes = bk.bytes_reduced - b0
        bk.time_comm = False
        t = torch.tensor([t_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        stg = bk.staging if bk.staging is not None else bk.grads
        for _ in range(2):
            bk.all_reduce(stg)
        torch.cuda.synchronize()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for _ in range(5):
            bk.all_reduce(stg)
        a1.record()
        torch.cuda.synchronize()
        alone = torch.tensor([a0.elapsed_time(a1) / 5], device=dev)
        dist.all_reduce(alone, op=dist.ReduceOp.MAX)
        sbytes = stg.numel() * stg.element_size()
        # (4) exposed communication: step time minus the time of the same step with the collective
        #     stubbed out (pack / unpack kernels still run), and minus the step with no gradient sync at all.
        #     Last, because the replicas drift apart once their gradients are no longer averaged.
        bk.stub_collective = True
        train_step(resident, False)
        ms_stub = timed(lambda: train_step(resident, False), args.steps)
        bk.stub_collective = False
        train_step(resident, False, sync=False)
        ms_nosync = timed(lambda: train_step(resident, False, sync=False), args.steps)
        step_ms = ms / args.steps
        comm = {"bytes_per_rank_per_step": nbytes, "ms_on_comm_stream": t.item(),
                "standalone_ms": alone.item(),
                "standalone_busbw_GBps": sbytes * 2 * (world - 1) / world / (alone.item() / 1e3) / 1e9,
                "busbw_GBps": nbytes * 2 * (world - 1) / world / (t.item() / 1e3) / 1e9,
                "nominal_GBps": 900.0, "bucket_dtype": args.bucket_dtype,
                "exposed_comm_ms": step_ms - ms_stub / args.steps,
                "exposed_comm_frac": (step_ms - ms_stub / args.steps) / step_ms,
                "step_ms_collective_stubbed": ms_stub / args.steps,
                "ddp_overhead_ms": step_ms - ms_nosync / args.steps,
                "step_ms_no_grad_sync": ms_nosync / args.steps,
                "comm": bk.describe()}

    torch_gpu = None
    if rank == 0 and world == 1 and not args.no_torch_baseline:
        del resident
        model = raw = opt = None
        torch.cuda.empty_cache()
        try:
            torch_gpu = torch_gpu_baseline(dev, wl)
        except Exception as exc:  # informational only
            torch_gpu = {"error": repr(exc)[:200]}
    if rank == 0:
        cpu = None
        if not args.no_cpu_baseline:
            tok_s, _, cores, sample, _ = cpu_reference_run(wl, world, 4, 1)
            cpu = {"value": tok_s, "unit": "tokens/s", "cores": cores, "kind": "port", "sample": sample}
        line = {
            "metric": "train_tokens_per_sec", "value": value, "unit": "tokens/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": bench_config(args.config, wl, world, accum, tokens_per_step, args.bucket_dtype),
            "mfu": {"flops_per_token": FLOPS_PER_TOKEN,
                    "vs_nominal_2.25PF": value * FLOPS_PER_TOKEN / (world * NOMINAL_BF16),
                    "vs_measured_peak": value * FLOPS_PER_TOKEN / (world * peaks["tf"] * 1e12),
                    "vs_measured_sustained": value * FLOPS_PER_TOKEN / (world * peaks["tf_sus"] * 1e12),
                    "measured_peak_tflops": peaks["tf"], "measured_sustained_tflops": peaks["tf_sus"]},
            "e2e": {"value": e2e_value, "unit": "tokens/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches), "host": host, "loss_at_init": loss0, "clocks": clocks, "roofline": roofline,
            "cpu_baseline": cpu, "torch_gpu_baseline": torch_gpu,
        }
        if comm is not None:
            line["allreduce"] = comm
            line["ddp_check"] = ddp_check
        emit(line)
    if ddp:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
