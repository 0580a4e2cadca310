"""Are NCCL's bf16 all-reduce results the same BITS on every rank?  (torchrun, one rank per GPU.)
Replicas stay identical only if they are; bench.py's ddp_check saw one rank drift in a 112 KB run of one bucket.
Each iteration: rank-dependent data, all-reduce on a side stream while GEMMs keep the GPU busy on the main
stream (as in the last micro-step's backward), then the result is compared with rank 0's copy."""
import os
import sys
import torch
import torch.distributed as dist

dist.init_process_group("nccl")
rank, world = dist.get_rank(), dist.get_world_size()
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
torch.cuda.set_device(dev)
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
sizes = [7_080_960, 1_769_472, 38_633_472]
side = torch.cuda.Stream()
a = torch.randn(4096, 4096, device=dev, dtype=torch.bfloat16)
b = torch.randn(4096, 4096, device=dev, dtype=torch.bfloat16)
g = torch.Generator(device=dev)
report = {}
for n in sizes:
    bad_iters, worst = 0, 0
    buf = torch.empty(n, device=dev, dtype=torch.bfloat16)
    ref = torch.empty_like(buf)
    for it in range(iters):
        g.manual_seed(1000 * it + rank)
        buf.copy_((torch.randn(n, device=dev, generator=g) * 0.01).to(torch.bfloat16))
        ev = torch.cuda.current_stream().record_event()
        for _ in range(6):
            c = a @ b                      # keeps the SMs busy while the collective runs
        side.wait_event(ev)
        with torch.cuda.stream(side):
            dist.all_reduce(buf)
        torch.cuda.current_stream().wait_stream(side)
        ref.copy_(buf)
        dist.broadcast(ref, src=0)
        ne = int((ref != buf).sum().item())
        t = torch.tensor([ne], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if t.item():
            bad_iters += 1
            worst = max(worst, int(t.item()))
            if ne:
                idx = (ref != buf).nonzero().flatten()
                print(f"rank {rank} size {n} iter {it}: {ne} elements differ from rank 0, span [{int(idx.min())}, {int(idx.max())}]",
                      file=sys.stderr, flush=True)
    report[n] = (bad_iters, worst)
if rank == 0:
    print({"world": world, "iters": iters, "NCCL_ALGO": os.environ.get("NCCL_ALGO", "(default)"),
           "mismatching_iterations / worst count per size": report})
dist.destroy_process_group()
