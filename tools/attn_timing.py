"""Phase timestamps of one CTA of the tcgen05 attention forward (clock64 deltas)."""
import torch
from nanosandbox_b200 import _lib, ops
lib = _lib.load()
B, T, H, D = 12, 1024, 12, 64
C = H * D
qkv = torch.randn(B, T, 3 * C, device="cuda").to(torch.bfloat16)
y = torch.empty(B, T, C, device="cuda", dtype=torch.bfloat16)
lse = torch.empty(B, H, T, device="cuda")
dbg = torch.zeros(1024 + 3 * 1152 + 16, dtype=torch.int64, device="cuda")
for _ in range(2):
    ops.attention_fwd(qkv, y, lse, B, T, H, D)
lib.nsb_attention_debug_timing(dbg.data_ptr())
ops.attention_fwd(qkv, y, lse, B, T, H, D)
torch.cuda.synchronize()
lib.nsb_attention_debug_timing(None)
d = dbg.cpu().tolist()
t0 = d[512]
print("softmax warp: per block [wait s_full start, s_full seen, softmax done, p_full arrived] (clocks since MMA start)")
for j in range(8):
    print(j, [x - t0 for x in d[4 * j:4 * j + 4]], " softmax clks", d[4 * j + 2] - d[4 * j + 1], " wait clks", d[4 * j + 1] - d[4 * j])
print("mma thread: per block [loop top, kv_full seen, S issued, p_full seen]")
for j in range(8):
    print(j, [x - t0 for x in d[512 + 4 * j:512 + 4 * j + 4]])

import collections
tr = dbg.cpu()[1024:1024 + 3 * 1152].view(1152, 3)
t_min = int(tr[:, 0].min()); t_max = int(tr[:, 1].max())
print("kernel span us", (t_max - t_min) / 1e3)
dur = (tr[:, 1] - tr[:, 0]).float() / 1e3
for qb in range(8):
    sel = dur[(7 - qb) * 144:(8 - qb) * 144]   # blockIdx.y = 7-qb
    st = (tr[(7 - qb) * 144:(8 - qb) * 144, 0] - t_min).float() / 1e3
    print(f"qb={qb} blocks={qb+1} dur us mean {sel.mean():.2f} min {sel.min():.2f} max {sel.max():.2f} | start us min {st.min():.2f} max {st.max():.2f}")
per_sm = collections.Counter(int(x) for x in tr[:, 2])
print("CTAs per SM: min", min(per_sm.values()), "max", max(per_sm.values()), "SMs used", len(per_sm))
busy = collections.defaultdict(float)
for i in range(1152):
    busy[int(tr[i, 2])] += float(dur[i])
print("per-SM sum of CTA durations us: min %.1f max %.1f" % (min(busy.values()), max(busy.values())))
