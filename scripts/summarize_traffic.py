#!/usr/bin/env python
"""Per-launch DRAM traffic of the GEMM launches of one score-network call, from
`ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv`: totals and a per-instantiation
table, written as JSON (bench.py reads `dram_bytes_per_launch` for `roofline.traffic`)."""
import csv
import json
import re
import sys
from collections import defaultdict

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3,
        "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}


def main(path, out):
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    per = defaultdict(dict)
    for r in csv.DictReader(lines):
        v = float(r["Metric Value"].replace(",", "")) * UNIT[r["Metric Unit"]]
        per[int(r["ID"])][r["Metric Name"]] = v
        per[int(r["ID"])]["name"] = re.sub(r"^void (ramp::)?|\(.*$", "", r["Kernel Name"])
    tot_r = sum(p["dram__bytes_read.sum"] for p in per.values())
    tot_w = sum(p["dram__bytes_write.sum"] for p in per.values())
    tot_ms = sum(p["gpu__time_duration.sum"] for p in per.values())
    by = defaultdict(lambda: dict(launches=0, read_GB=0.0, write_GB=0.0, ncu_ms=0.0))
    for p in per.values():
        b = by[p["name"]]
        b["launches"] += 1
        b["read_GB"] += p["dram__bytes_read.sum"] / 1e9
        b["write_GB"] += p["dram__bytes_write.sum"] / 1e9
        b["ncu_ms"] += p["gpu__time_duration.sum"]
    for b in by.values():
        b["GBps"] = (b["read_GB"] + b["write_GB"]) / (b["ncu_ms"] * 1e-3)
    doc = {"source": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none "
                     "--kernel-name regex:gemm_tc_kernel --launch-skip 260 --launch-count 260 python scripts/profile_unet.py "
                     "(all 260 GEMM launches of one score-network call on 7168 rows)",
           "launches": len(per), "dram_bytes_read": tot_r, "dram_bytes_write": tot_w,
           "dram_bytes_per_launch": (tot_r + tot_w) / max(1, len(per)), "ncu_time_ms": tot_ms,
           "by_instantiation": dict(sorted(by.items()))}
    with open(out, "w") as f:
        json.dump(doc, f, indent=1)
    print("launches %d  read %.1f GB  write %.1f GB  ncu %.2f ms" % (len(per), tot_r / 1e9, tot_w / 1e9, tot_ms))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
