#!/usr/bin/env python
"""Per-source-line stall samples / executed instructions of one captured kernel, from an .ncu-rep taken with
--import-source on:   python tools/ncu_lines.py rep.ncu-rep [kernel-substring] [top]"""
import collections
import csv
import io
import subprocess
import sys

def num(x):
    try:
        return int(float(x))
    except (ValueError, TypeError):
        return 0


rep = sys.argv[1]
pat = sys.argv[2] if len(sys.argv) > 2 else ""
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
agg = collections.defaultdict(lambda: [0, 0, collections.Counter(), ""])
fpath, fn, head, seen_fn = None, None, None, set()
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fpath = r[1].split("/")[-1]
    elif r[0] == "Function Name":
        fn = r[1]
    elif r[0] == "Line No":
        head = r
        cS, cI = head.index("# Samples"), head.index("Instructions Executed")
        stall = [(i, h) for i, h in enumerate(head) if h.startswith("stall_") and "Not Issued" not in h]
    elif head and fn and pat in fn and len(r) > cI and r[0].isdigit():
        if fn not in seen_fn:
            if seen_fn:
                continue            # first matching launch only
            seen_fn.add(fn)
        a = agg[(fpath, int(r[0]))]
        a[0] += num(r[cS])
        a[1] += num(r[cI])
        a[3] = r[1].strip()[:100]
        for i, h in stall:
            v = num(r[i])
            if v:
                a[2][h] += v
ts, ti = sum(a[0] for a in agg.values()), sum(a[1] for a in agg.values())
print("kernel:", list(seen_fn)[:1], "samples", ts, "warp-instructions", ti)
print("--- top lines by stall samples")
for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{f}:{ln:4d} s={a[0]:5d} {100 * a[0] / max(ts, 1):4.1f}% i={100 * a[1] / max(ti, 1):4.1f}% {dict(a[2].most_common(2))} | {a[3]}")
print("--- top lines by executed instructions")
for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{f}:{ln:4d} i={a[1]:9d} {100 * a[1] / max(ti, 1):4.1f}% s={100 * a[0] / max(ts, 1):4.1f}% | {a[3]}")
