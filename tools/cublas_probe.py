"""What cuBLAS launches for the per-layer shapes (run under ncu on the GPU box; informational)."""
import torch
bf = torch.bfloat16
shapes = [(12288, 2304, 768), (12288, 768, 768), (12288, 3072, 768), (12288, 768, 3072)]
for M, N, K in shapes:
    a = torch.randn(M, K, device="cuda").to(bf); b = torch.randn(N, K, device="cuda").to(bf)
    for _ in range(3):
        c = a @ b.t()
    torch.cuda.synchronize()
# wgrad form: dY^T X
a = torch.randn(12288, 2304, device="cuda").to(bf); b = torch.randn(12288, 768, device="cuda").to(bf)
for _ in range(3):
    c = a.t() @ b
torch.cuda.synchronize()
print("done")
