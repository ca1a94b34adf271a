#!/usr/bin/env python
"""Profiling driver for the nearest-point kernel at the sweep's shape (8192 x 64 waypoints, 1 M cloud points)."""
import sys
import torch

sys.path.insert(0, ".")
from ramp_b200 import ops  # noqa: E402

dim = int(sys.argv[1]) if len(sys.argv) > 1 else 2
dev = torch.device("cuda", 0)
g = torch.Generator("cpu").manual_seed(7)
P, Bq, H = 1 << 20, 8192, 64
centres = torch.rand(256, dim, generator=g) * 2 - 1
pts = (centres[torch.randint(0, 256, (P,), generator=g)] + 0.05 * torch.randn(P, dim, generator=g)).to(dev)
x = (torch.rand(Bq, H, 2 * dim, generator=g) * 2 - 1).to(dev)
cloud = ops.CloudIndex(pts)
for _ in range(3):
    d, i = cloud.query(x, 0.07)
torch.cuda.synchronize()
print("ok", float(d[torch.isfinite(d)].mean()))
