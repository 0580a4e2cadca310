"""Per-kernel SASS evidence from the built library (no GPU needed):

    python tools/sass_summary.py [--excerpt 40] > profiles/rNN_sass_summary.txt

For every kernel in nanosandbox_b200/lib/libnanosandbox_b200.so: instruction count and the number of
Blackwell-specific instructions (B200_PROFILING.md, "What proves a Blackwell-native kernel"):
UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG/UTMAREDG = TMA tensor copies,
UBLKCP = cp.async.bulk, SYNCS = mbarrier ops, HMMA = legacy mma.sync.  With --excerpt N the N
instructions around the first UTC*MMA (the mainloop) of each tensor-core kernel are listed too.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "nanosandbox_b200", "lib", "libnanosandbox_b200.so")
PATTERNS = [("UTCMMA", r"\bUTC[A-Z]*MMA"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"), ("UTMALDG", r"\bUTMALDG"),
            ("UTMASTG", r"\bUTMASTG"), ("UTMAREDG", r"\bUTMAREDG"), ("UBLKCP", r"\bUBLKCP"), ("SYNCS", r"\bSYNCS"),
            ("HMMA", r"\bHMMA"), ("MUFU", r"\bMUFU"), ("ATOMG/REDG", r"\b(ATOMG|REDG|RED)\b"), ("UTCBAR.MULTICAST", r"UTCBAR[.A-Z0-9_]*MULTICAST"), ("TMA-load multicast", r"(UTMALDG|UBLKCP)[.A-Z0-9_]*MULTICAST")]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def main():
    excerpt = int(sys.argv[sys.argv.index("--excerpt") + 1]) if "--excerpt" in sys.argv else 0
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in sass.split("\n"):
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = []
            continue
        if cur is not None and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
            kernels[cur].append(line.rstrip())
    names = demangle(list(kernels))
    print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)} ({os.path.getsize(LIB)} bytes), {len(kernels)} kernels, sm_100a")
    print("# " + " ".join(f"{n:>10s}" for n, _ in PATTERNS) + "     instrs  kernel")
    listed = []
    for k, lines in kernels.items():
        text = "\n".join(lines)
        counts = [len(re.findall(p, text)) for _, p in PATTERNS]
        short = re.sub(r"\(.*", "", names[k])[:110]
        print("  " + " ".join(f"{c:10d}" for c in counts) + f" {len(lines):10d}  {short}")
        if counts[0] and excerpt:
            listed.append((short, lines))
    if excerpt:
        seen = set()
        for short, lines in listed:
            base = re.sub(r"<.*", "", short)
            if base in seen:          # one excerpt per kernel template
                continue
            seen.add(base)
            i = next(j for j, l in enumerate(lines) if re.search(PATTERNS[0][1], l))
            print(f"\n## {short}: {excerpt} instructions around the first tcgen05.mma")
            for l in lines[max(0, i - excerpt // 2): i + excerpt // 2]:
                print(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/", "", l))


if __name__ == "__main__":
    main()
