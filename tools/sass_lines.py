#!/usr/bin/env python
"""Static SASS instruction count per source line of one kernel (nvdisasm -g on the in-tree object): where a kernel's
instruction budget goes before spending GPU time.   python tools/sass_lines.py <object.o> <kernel-name-substring> [min]"""
import collections
import os
import re
import subprocess
import sys
import tempfile

obj, pat = sys.argv[1], sys.argv[2]
thr = int(sys.argv[3]) if len(sys.argv) > 3 else 20
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, check=True, capture_output=True)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout
cnt, cur, f, line, infn = collections.Counter(), None, None, None, False
for l in txt.split("\n"):
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
    if m:
        infn = pat in m.group(1)
        continue
    if l.lstrip().startswith(".section"):
        infn = False
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        f, line = os.path.basename(m.group(1)), int(m.group(2))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l):
        cnt[(f, line)] += 1
print("total", sum(cnt.values()))
byfile = collections.Counter()
for (f, ln), c in cnt.items():
    byfile[f] += c
print(dict(byfile))
srcs = {}
for (f, ln), c in sorted(cnt.items(), key=lambda kv: (kv[0][0] or "", kv[0][1] or 0)):
    if c >= thr:
        if f not in srcs:
            p = os.path.join(os.path.dirname(os.path.abspath(obj)), "..", f)
            srcs[f] = open(p).read().split("\n") if os.path.exists(p) else None
        s = srcs[f][ln - 1].strip()[:100] if srcs[f] else ""
        print(f"{f}:{ln:5d} {c:5d}  {s}")
