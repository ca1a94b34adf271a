#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel: launches, total / mean time, share."""
import csv
import re
import sys
from collections import defaultdict


def main(path):
    rows = []
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((r["Kernel Name"], float(r["Metric Value"].replace(",", "")), r["Grid Size"], r["Block Size"]))
    agg = defaultdict(lambda: [0, 0.0])
    for name, ns, *_ in rows:
        short = re.sub(r"\(.*", "", name)
        agg[short][0] += 1
        agg[short][1] += ns
    total = sum(v[1] for v in agg.values())
    print("# %s: %d launches, %.3f ms total (per-launch times are cold-cache, serialised; compare SHARES)" % (path, len(rows), total / 1e6))
    print("%-46s %8s %12s %10s %7s" % ("kernel", "launches", "total_ms", "mean_us", "share"))
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-46s %8d %12.3f %10.2f %6.1f%%" % (k, n, ns / 1e6, ns / 1e3 / n, 100 * ns / total))


if __name__ == "__main__":
    main(sys.argv[1])
