"""Summarise an ncu --csv launch list (gpu__time_duration.sum): per-kernel count, total, share, average."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hdr = next(r for r in rows if "Kernel Name" in r)
start = rows.index(hdr)
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
ui = hdr.index("Metric Unit") if "Metric Unit" in hdr else None
agg = collections.OrderedDict()
for r in rows[start + 1:]:
    if len(r) <= vi:
        continue
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    if ui is not None and r[ui] in ("ns", "nsecond"):
        v /= 1e3
    elif ui is not None and r[ui] in ("ms", "msecond"):
        v *= 1e3
    name = r[ki].split("(")[0][:88]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
if len(sys.argv) > 2:
    print(sys.argv[2])
for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{name:88s} n={n:4d} total={us:10.1f} us share={us / tot * 100:5.1f}% avg={us / n:8.1f} us")
print(f"total us {tot:.1f}")
