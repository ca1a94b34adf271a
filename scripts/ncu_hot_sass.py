#!/usr/bin/env python
"""Top SASS instructions by stall samples from `ncu -i X.ncu-rep --page source --csv --print-source sass`."""
import csv
import sys

path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
for n, r in enumerate(rows[hi + 1:]):
    if len(r) < len(hdr):
        continue
    try:
        s = int(r[col["# Samples"]] or 0)
    except ValueError:
        continue
    data.append((n, s, r))
tot = sum(s for _, s, _ in data)
print("total samples", tot, "instructions", len(data))
for n, s, r in sorted(data, key=lambda t: -t[1])[:top]:
    st = sorted(((int(r[col[h]] or 0), h[6:]) for h in stalls), reverse=True)[:2]
    print("%5d %5.1f%% ex=%-9s %-70s %s" % (n, 100.0 * s / tot, r[col["Instructions Executed"]], r[col["Source"]][:70],
                                          " ".join("%s:%d" % (h, v) for v, h in st if v)))
