#!/usr/bin/env python
"""Compare two `bench.py --dump-launches` CSVs per (class, tag): total ms old -> new."""
import csv
import sys
from collections import defaultdict


def load(p):
    a = defaultdict(float)
    for r in csv.DictReader(open(p)):
        a[(r["class"], r["tag"])] += float(r["ms"])
    return a


a, b = load(sys.argv[1]), load(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
print("total %.1f -> %.1f" % (sum(a.values()), sum(b.values())))
mode = defaultdict(lambda: [0.0, 0.0])
for k in set(a) | set(b):
    m = k[1].rsplit("_e", 1)[-1] if k[0] == "gemm" else k[0]
    mode[m][0] += a.get(k, 0); mode[m][1] += b.get(k, 0)
for m, (x, y) in sorted(mode.items()):
    print("  mode %-10s %8.1f -> %8.1f" % (m, x, y))
for k in sorted(set(a) | set(b), key=lambda k: -max(a.get(k, 0), b.get(k, 0)))[:top]:
    print("%-10s %-34s %8.2f -> %8.2f" % (k[0], k[1], a.get(k, 0), b.get(k, 0)))
