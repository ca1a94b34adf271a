#!/usr/bin/env python
"""Sweep the nearest-point kernel's (R, TILE) choices at the sweep's shape in ONE process (RAMP_NN_FORCE is read per call).
Prints ms, fraction of the nominal FP32 peak and the per-block cost  ms / (ceil(CTAs / SMs) * R)  that csrc/field_ops.cu's
nn_choose tables hold.  usage: sweep_field.py [Bq]"""
import os
import sys
import torch

sys.path.insert(0, ".")
from ramp_b200 import ops  # noqa: E402

Bq = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
P_LOG2 = int(sys.argv[2]) if len(sys.argv) > 2 else 20
dev = torch.device("cuda", 0)
P, H = 1 << P_LOG2, 64
sms = torch.cuda.get_device_properties(0).multi_processor_count
peak = sms * 128 * 2 * 1.965e9
for dim in (2, 3):
    g = torch.Generator("cpu").manual_seed(7)
    centres = torch.rand(256, dim, generator=g) * 2 - 1
    pts = (centres[torch.randint(0, 256, (P,), generator=g)] + 0.05 * torch.randn(P, dim, generator=g)).to(dev)
    x = (torch.rand(Bq, H, 2 * dim, generator=g) * 2 - 1).to(dev)
    cloud = ops.CloudIndex(pts)
    d = torch.empty(Bq, H, device=dev)
    i = torch.empty(Bq, H, device=dev, dtype=torch.int32)
    ref = None
    for force in [None] + [(r, 512) for r in (1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 14, 16)]:
        if force is None:
            os.environ.pop("RAMP_NN_FORCE", None)
        else:
            os.environ["RAMP_NN_FORCE"] = "%d,%d" % force
        for _ in range(2):
            cloud.query(x, 0.07, d, i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(3):
            cloud.query(x, 0.07, d, i)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        chk = int(i.long().sum())
        if ref is None:
            ref = chk
        line = "dim %d %-10s %7.2f ms  frac %.3f  checksum %s" % (
            dim, "auto" if force is None else "R=%d,T=%d" % force, ms, Bq * H * P * (3 * dim - 1) / (ms * 1e-3) / peak,
            "ok" if chk == ref else "MISMATCH %d vs %d" % (chk, ref))
        if force is not None:
            r = force[0]
            ctas = (Bq * H + 128 * r - 1) // (128 * r)
            line += "  cost/block %.3f" % (ms / (((ctas + sms - 1) // sms) * r))
        print(line, flush=True)
os.environ.pop("RAMP_NN_FORCE", None)
