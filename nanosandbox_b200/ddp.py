"""Data-parallel wrapper: our own gradient bucket manager in place of
``torch.nn.parallel.DistributedDataParallel`` + the C++ Reducer
(torch/nn/parallel/distributed.py:661-674, 801, 881-890; torch/include/torch/csrc/distributed/c10d/
reducer.hpp) on the path ``train.py`` takes when RANK is set (/root/reference/README.md:6-8).

Same contract as DDP for nanoGPT's loop: ``.module``,
``.require_backward_grad_sync`` written every micro-step, ``__call__`` -> ``(logits, loss)``,
rank-0 parameters broadcast at construction, and after the last micro-step's ``backward()``
every ``.grad`` is the mean over ranks of the locally accumulated gradient.

Differences by design: a bucket is a contiguous slice of the flat gradient arena (no copies in or
out of bucket views); the 1/world scaling is fused into the pack kernel that casts the slice to
bf16; the all-reduce runs over NCCL (NVLink 5 / NVSwitch) on a side stream as soon as backward has
finalised that slice, overlapping the rest of backward.  ``bucket_dtype='fp32'`` keeps exact
DDP numerics (scale in place, all-reduce the fp32 slice itself).
"""
from __future__ import annotations

import contextlib
import os
from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist


class GradBucketManager:
    """Host logic of the bucketed all-reduce; kernels are injected so that the logic can be driven
    on CPU/gloo in tests while the product path uses the CUDA pack/unpack kernels."""

    def __init__(self, grads: torch.Tensor, ranges: List[Tuple[int, int]], world_size: int,
                 process_group=None, bucket_dtype: str = "bf16",
                 pack: Optional[Callable] = None, unpack: Optional[Callable] = None,
                 scale: Optional[Callable] = None):
        assert bucket_dtype in ("bf16", "fp32")
        self.grads, self.ranges, self.world = grads, ranges, world_size
        self.pg = process_group
        self.bucket_dtype = bucket_dtype
        self.pack, self.unpack, self.scale = pack, unpack, scale
        self.cuda = grads.is_cuda
        self.staging = (torch.empty(grads.numel(), dtype=torch.bfloat16, device=grads.device)
                        if bucket_dtype == "bf16" else None)
        self.comm_stream = torch.cuda.Stream(device=grads.device, priority=-1) if self.cuda else None
        self.launched: List[int] = []
        self.bytes_reduced = 0
        self.comm_events: List[Tuple[object, object]] = []
        self.time_comm = False
        self.stub_collective = False    # measurement hook: skip the collective itself (pack / unpack still run)
        self.comm_info = {}
        # NSB_DDP_VERIFY=1: after every synchronised backward, check that the gradient arena is the same bits on
        # every rank (one checksum per bucket, one small all-reduce) and name the first bucket that is not: replicas
        # stay identical only if it is, and the step at which they stop says what ran beside it
        # bf16 buckets are written back a second time, on the compute stream, once backward has ended (see finish)
        self.reunpack = os.environ.get("NSB_DDP_REUNPACK", "1") != "0"
        self.verify = os.environ.get("NSB_DDP_VERIFY", "0") == "1"
        self.verify_steps = 0
        self.verify_failures: List[dict] = []

    def all_reduce(self, t: torch.Tensor) -> None:
        if not self.stub_collective:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.pg)

    def describe(self) -> dict:
        return dict(self.comm_info, buckets=len(self.ranges), bucket_dtype=self.bucket_dtype)

    def reduce_bucket(self, k: int) -> None:
        s, e = self.ranges[k]
        g = self.grads[s:e]
        inv = 1.0 / self.world
        if self.cuda:
            ev = torch.cuda.current_stream().record_event()
            self.comm_stream.wait_event(ev)
            ctx = torch.cuda.stream(self.comm_stream)
        else:
            ctx = contextlib.nullcontext()
        with ctx:
            t0 = t1 = None
            if self.cuda and self.time_comm:
                t0 = torch.cuda.Event(enable_timing=True)
                t1 = torch.cuda.Event(enable_timing=True)
                t0.record()
            if self.bucket_dtype == "bf16":
                st = self.staging[s:e]
                self.pack(g, st, inv)                       # cast fp32 -> bf16 and scale by 1/world
                self.all_reduce(st)
                self.unpack(st, g)                          # bf16 -> fp32 back into the arena slice
                self.bytes_reduced += st.numel() * 2
            else:
                self.scale(g, inv)                          # divide first, then sum: DDP's order
                self.all_reduce(g)
                self.bytes_reduced += g.numel() * 4
            if t0 is not None:
                t1.record()
                self.comm_events.append((t0, t1))
        self.launched.append(k)

    def finish(self) -> None:
        """Fence: the compute stream waits for every bucket launched during this backward."""
        if self.cuda:
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        assert self.launched == list(range(len(self.ranges))), \
            f"buckets reduced out of order or missing: {self.launched}"
        self.launched = []
        if self.cuda and self.reunpack and self.bucket_dtype == "bf16" and self.unpack is not None:
            # Belt and braces for the replica drift seen on 8 GPUs: whatever reached a bucket's
            # slice of the arena after its first write-back - the signature was the tail of a weight-gradient GEMM's
            # accumulation landing while the bucket was being unpacked - is overwritten from the staging buffer,
            # which nothing but the pack kernel and NCCL write.  Runs behind every backward kernel of the compute
            # stream; 0.75 GB of traffic, ~0.1 ms per optimizer step.
            for s, e in self.ranges:
                self.unpack(self.staging[s:e], self.grads[s:e])
        if self.verify and not self.stub_collective:
            self.verify_replicas()

    def bucket_checksums(self) -> torch.Tensor:
        """One int64 per bucket: the sum of the gradient words read as integers (any differing bit changes it,
        up to cancellation)."""
        words = self.grads.view(torch.int32)
        return torch.stack([words[s:e].sum(dtype=torch.int64) for s, e in self.ranges])

    def verify_replicas(self) -> bool:
        """True when every rank holds the same gradient bits; otherwise records {step, buckets} in
        verify_failures (and says so on stderr).  Costs one pass over the gradients: a debugging aid."""
        mine = self.bucket_checksums()
        lo, hi = mine.clone(), mine.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=self.pg)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=self.pg)
        self.verify_steps += 1
        bad = (lo != hi).nonzero().flatten().tolist()
        if bad:
            rec = {"step": self.verify_steps, "buckets": bad,
                   "this_rank_differs_from_min": [k for k in bad if int(mine[k]) != int(lo[k])]}
            self.verify_failures.append(rec)
            import sys
            print(f"nanosandbox_b200.ddp: synchronised gradients differ between ranks: {rec}", file=sys.stderr, flush=True)
        return not bad


def gradient_communicator(max_ctas: int):
    """A communicator of its own for the gradient all-reduce, with a bound on the CTAs NCCL may put on the
    GPU (ncclConfig_t.maxCTAs).  The buckets are reduced while the last micro-step's backward is still
    running; every SM an NCCL CTA occupies is an SM a GEMM CTA cannot use, and over NVLink 5 / NVSwitch a
    few CTAs already move 250 MB in about a millisecond, far inside the ~14 ms of backward it hides under.
    Returns (process_group, info)."""
    info = {"communicator": "default process group", "max_ctas": None}
    if max_ctas <= 0 or dist.get_backend() != "nccl":
        return None, info
    try:
        from torch.distributed import ProcessGroupNCCL
        opts = ProcessGroupNCCL.Options()
        opts.is_high_priority_stream = True
        opts.config.max_ctas = int(max_ctas)
        opts.config.min_ctas = 1
        pg = dist.new_group(ranks=list(range(dist.get_world_size())), backend="nccl", pg_options=opts)
        info = {"communicator": "dedicated NCCL communicator (high-priority stream)", "max_ctas": int(max_ctas)}
        return pg, info
    except Exception as exc:  # an NCCL / torch build without the config field: keep the default group
        info["note"] = f"dedicated communicator unavailable: {type(exc).__name__}: {exc}"[:200]
        return None, info


class DistributedDataParallel:
    def __init__(self, module, device_ids=None, process_group=None, bucket_dtype: str = "bf16",
                 broadcast_parameters: bool = True, max_ctas: Optional[int] = None):
        from . import ops
        self.module = module
        self.require_backward_grad_sync = True
        self.world_size = dist.get_world_size(process_group)
        comm_info = {"communicator": "caller's process group", "max_ctas": None}
        if process_group is None:
            if max_ctas is None:
                # measured on 8 x B200 (profiles/r02_scale/): a cap of 8 CTAs and NCCL's own choice give the same
                # step time within noise (71.0 vs 70.8 ms), so the default is NCCL's; the cap stays available
                max_ctas = int(os.environ.get("NSB_NCCL_MAX_CTAS", "0"))
            grad_pg, comm_info = gradient_communicator(max_ctas)
        else:
            grad_pg = process_group
        self.process_group = process_group
        arena = module.arena()
        if broadcast_parameters:
            # one broadcast over the flat parameter arena (DDP: _sync_module_states, 250 MiB chunks)
            dist.broadcast(arena.params, src=0, group=process_group)
            arena.lowp_version = -1
        self.buckets = GradBucketManager(
            arena.grads, module.bucket_ranges, self.world_size, grad_pg, bucket_dtype,
            pack=lambda g, st, inv: ops.cast_f32_to_bf16(g, st, inv),
            unpack=lambda st, g: ops.cast_bf16_to_f32(st, g),
            scale=lambda g, inv: ops.scale_f32(g, inv))
        self.buckets.comm_info = comm_info
        module._grad_sync = self

    # hooks called by GPT._engine_backward on the compute stream
    def bucket_ready(self, k: int) -> None:
        if self.require_backward_grad_sync:
            self.buckets.reduce_bucket(k)

    def finish_backward(self) -> None:
        if self.require_backward_grad_sync:
            self.buckets.finish()

    @contextlib.contextmanager
    def no_sync(self):
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    def __call__(self, *a, **k):
        return self.module(*a, **k)

    forward = __call__

    def parameters(self, *a, **k):
        return self.module.parameters(*a, **k)

    def named_parameters(self, *a, **k):
        return self.module.named_parameters(*a, **k)

    def state_dict(self, *a, **k):
        return {"module." + kk: v for kk, v in self.module.state_dict(*a, **k).items()}

    def train(self, mode=True):
        self.module.train(mode)
        return self

    def eval(self):
        self.module.eval()
        return self
