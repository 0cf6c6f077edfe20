#!/usr/bin/env python
"""Per-kernel time of ONE bench step from an ncu launch list of `bench.py --steps 1 --warmup 1 --only-value` (three identical
steps: the last third of every kernel's launches is taken):  python profiles/tools/kernel_times.py launches.csv"""
import collections
import csv
import sys
U = {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}
rows = [r for r in csv.DictReader(l for l in open(sys.argv[1]) if l.startswith('"'))]
t = collections.OrderedDict()
for r in rows:
    if r["Metric Name"] == "gpu__time_duration.sum":
        t.setdefault(r["Kernel Name"].split("(")[0].replace("void ", "").replace("pt::", ""), []).append(float(r["Metric Value"].replace(",", "")) * U[r["Metric Unit"]])
tot = 0.0
for n, v in sorted(t.items(), key=lambda kv: -sum(kv[1][-(len(kv[1]) // 3):])):
    k = len(v) // 3
    if k:
        tot += sum(v[-k:])
        print("%-32s %3d launches %8.3f ms" % (n, k, sum(v[-k:])))
print("%-32s %16.3f ms" % ("sum", tot))
