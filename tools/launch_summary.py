#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file X ...`):
    python tools/launch_summary.py X.csv "command line that was profiled" > X_summary.txt
Per-launch times under ncu are cold-cache and serialised: read the SHARES, not the absolutes."""
import collections
import csv
import re
import sys


def main(path, cmd=""):
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 10]
    h = rows[0]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}
    tot = collections.OrderedDict()
    for r in rows[1:]:
        name = re.sub(r"\(.*", "", r[ki]).replace("void ", "")
        name = re.sub(r"<unnamed>::", "", name)
        us = float(r[vi].replace(",", "")) * scale.get(r[ui], 1.0)
        t = tot.setdefault(name, [0, 0.0])
        t[0] += 1
        t[1] += us
    total = sum(t[1] for t in tot.values())
    n = sum(t[0] for t in tot.values())
    if cmd:
        print(cmd)
    print("per-launch times cold-cache and serialised (ncu); raw list: %s" % path)
    print("total %.0f us over %d launches" % (total, n))
    for name, (c, us) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        print("%-72s %5d launches %10.0f us %5.1f%% %8.1f us/launch" % (name[:72], c, us, 100 * us / total, us / c))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
