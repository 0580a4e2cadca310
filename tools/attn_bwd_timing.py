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
for _ in range(2):
    ops.attention_bwd(qkv, y, dy, lse, dqkv, delta, B, T, H, D, dq_acc)
lib.nsb_attention_debug_timing(dbg.data_ptr())
ops.attention_bwd(qkv, y, dy, lse, dqkv, delta, B, T, H, D, dq_acc)
torch.cuda.synchronize()
lib.nsb_attention_debug_timing(None)
d = dbg.cpu().tolist()
t0 = d[0]
print("compute warp per iteration: [top, sdp_full seen, P/dS in regs, dq_full(it-1) seen, stored+arrived, dQ written back]")
for it in range(8):
    r = [x - t0 for x in d[8 * it:8 * it + 6]]
    print(it, r, "wait_sdp", r[1] - r[0], "compute", r[2] - r[1], "wait_dq", r[3] - r[2], "store", r[4] - r[3], "writeback", r[5] - r[4])
print("inside the probabilities step (rel. sdp_full seen): [tmem 0 landed, math 0 done, tmem 1 landed, math 1 done]")
for it in range(8):
    print(it, [x - d[8 * it + 1] for x in d[256 + 4 * it:256 + 4 * it + 4]])
print("mma warp per iteration: [top, pds_full seen, sdp_free seen (before S/dP of the next block is issued), dq_free seen]")
for it in range(8):
    print(it, [x - t0 for x in d[512 + 4 * it:512 + 4 * it + 4]])

import collections
tr = dbg.cpu()[1024:1024 + 3 * 1152].view(1152, 3)
t_min = int(tr[:, 0].min()); t_max = int(tr[:, 1].max())
print("kernel span us", (t_max - t_min) / 1e3)
dur = (tr[:, 1] - tr[:, 0]).float() / 1e3
for jb in range(8):
    sel = dur[jb * 144:(jb + 1) * 144]
    st = (tr[jb * 144:(jb + 1) * 144, 0] - t_min).float() / 1e3
    print(f"key block {jb}: {8 - jb} query blocks, CTA duration us mean {sel.mean():.2f} min {sel.min():.2f} max {sel.max():.2f} | start us min {st.min():.2f} max {st.max():.2f}")
busy = collections.defaultdict(float)
for i in range(1152):
    busy[int(tr[i, 2])] += float(dur[i])
print("per-SM sum of CTA durations us: min %.1f max %.1f mean %.1f" % (min(busy.values()), max(busy.values()), sum(busy.values()) / len(busy)))
