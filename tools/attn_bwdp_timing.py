"""clock64 trace of CTA 0 of the persistent attention backward (attn_bwd_tc16p_kernel<true>) at the cfg4 shape:
one probability warp, the MMA warp and one dQ warp, per query block g of the CTA's walk."""
import torch
from nanosandbox_b200 import _lib, ops
lib = _lib.load()
B, T, H, D = 12, 1024, 12, 64
C = H * D
qkv = torch.randn(B, T, 3 * C, device="cuda").to(torch.bfloat16)
y = torch.empty(B, T, C, device="cuda", dtype=torch.bfloat16); dy = torch.randn_like(y)
lse = torch.empty(B, H, T, device="cuda"); dqkv = torch.empty_like(qkv)
delta = torch.empty(B, H, T, device="cuda"); dq_acc = torch.zeros(B * T, C, device="cuda")
dbg = torch.zeros(1024 + 3 * 1152 + 16, dtype=torch.int64, device="cuda")
ops.attention_fwd(qkv, y, lse, B, T, H, D)
for _ in range(3):
    ops.attention_bwd(qkv, y, dy, lse, dqkv, delta, B, T, H, D, dq_acc)
lib.nsb_attention_debug_timing(dbg.data_ptr())
ops.attention_bwd(qkv, y, dy, lse, dqkv, delta, B, T, H, D, dq_acc)
torch.cuda.synchronize()
lib.nsb_attention_debug_timing(None)
d = dbg.cpu().tolist()
t0 = d[0]
n = 36
print("probability warp, per block: top | stats barrier | wait S/dP | half 0 | half 1 | wait dq_full(g-1) | stores+fence+arrive || period")
prev = None
for g in range(n):
    r = d[8 * g:8 * g + 7]
    if r[0] == 0: break
    per = (r[0] - prev) if prev else 0
    prev = r[0]
    print(f"{g:2d} t={r[0] - t0:7d} | {r[1] - r[0]:5d} | {r[2] - r[1]:5d} | {r[3] - r[2]:5d} | {r[4] - r[3]:5d} | {r[5] - r[4]:5d} | {r[6] - r[5]:5d} || {per}")
print("probability warp on the diagonal (warp 4), same columns")
prev = None
for g in range(n):
    r = d[1536 + 8 * g:1536 + 8 * g + 7]
    if r[0] == 0: break
    per = (r[0] - prev) if prev else 0
    prev = r[0]
    print(f"{g:2d} t={r[0] - t0:7d} | {r[1] - r[0]:5d} | {r[2] - r[1]:5d} | {r[3] - r[2]:5d} | {r[4] - r[3]:5d} | {r[5] - r[4]:5d} | {r[6] - r[5]:5d} || {per}")
print("MMA warp, per block: top | wait sdp_free | wait Q/dO (K/V) | issue S/dP(next) | wait pds_full | wait dkv_free | issue dV dK | wait dq_free || period")
prev = None
for g in range(n):
    r = d[512 + 8 * g:512 + 8 * g + 8]
    if r[0] == 0: break
    per = (r[0] - prev) if prev else 0
    prev = r[0]
    if r[1] == 0:   # last block: no next
        print(f"{g:2d} t={r[0] - t0:7d} | (last) | wait pds {r[4] - r[3]:5d} | {r[5] - r[4]:5d} | {r[6] - r[5]:5d} | {r[7] - r[6]:5d} || {per}")
        continue
    print(f"{g:2d} t={r[0] - t0:7d} | {r[1] - r[0]:5d} | {r[2] - r[1]:5d} | {r[3] - r[2]:5d} | {r[4] - r[3]:5d} | {r[5] - r[4]:5d} | {r[6] - r[5]:5d} | {r[7] - r[6]:5d} || {per}")
print("dQ warp, per block: wait dq_full | read + first half staged (dq_free) | second half + TMA issue || period")
prev = None
for g in range(n):
    r = d[1024 + 4 * g:1024 + 4 * g + 4]
    if r[0] == 0: break
    per = (r[0] - prev) if prev else 0
    prev = r[0]
    print(f"{g:2d} t={r[0] - t0:7d} | {r[1] - r[0]:5d} | {r[2] - r[1]:5d} | {r[3] - r[2]:5d} || {per}")
