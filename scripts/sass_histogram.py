"""Per-kernel SASS instruction histogram of the shipped library: the mnemonics that prove which hardware path a kernel
uses (B200_PROFILING.md): UTCHMMA = tcgen05.mma, UTMALDG = TMA tensor load, UBLKCP = cp.async.bulk, LDTM / STTM =
tcgen05.ld / st, HMMA = legacy mma.sync, FFMA / MUFU for the CUDA-core kernels.
    python scripts/sass_histogram.py [lib.so] > profiles/rN_sass_histogram.txt        (no GPU needed: cuobjdump)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "ramp_b200", "libramp_b200.so")
KEYS = ["UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "UBLKCP", "LDTM", "STTM", "UTCBAR", "SYNCS", "HMMA", "LDSM", "LDGSTS",
        "FFMA2", "FFMA", "FMNMX3", "MUFU", "F2FP", "LDG", "STG", "LDS", "STS", "BAR"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()  # noqa: E731
    kernels, cur = collections.OrderedDict(), None
    for ln in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", ln)
        if m and cur is not None:
            cur[m.group(1)] += 1
            cur["_total"] += 1
    print("# %s  (cuobjdump -sass, sm_100a)" % os.path.relpath(LIB, ROOT))
    print("# counts of static SASS instructions per kernel; columns:", " ".join(KEYS))
    for name, c in kernels.items():
        short = re.sub(r"\(.*", "", demangle(name))
        short = short.replace("void ramp::", "").replace("ramp::", "")
        cells = " ".join("%s=%d" % (k, c[k]) for k in KEYS if c[k])
        print("%-44s total=%-6d %s" % (short[:44], c["_total"], cells))
    tot = collections.Counter()
    for c in kernels.values():
        tot.update(c)
    print("# library totals:", " ".join("%s=%d" % (k, tot[k]) for k in KEYS if tot[k]))


if __name__ == "__main__":
    main()
