#!/usr/bin/env python
"""Time the nearest-point kernel at the sweep's shape: time_field.py DIM  (honours RAMP_NN_FORCE=R,TILE)."""
import sys
import torch

sys.path.insert(0, ".")
from ramp_b200 import ops  # noqa: E402

dim = int(sys.argv[1]) if len(sys.argv) > 1 else 2
Bq = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
dev = torch.device("cuda", 0)
g = torch.Generator("cpu").manual_seed(7)
P, H = 1 << 20, 64
centres = torch.rand(256, dim, generator=g) * 2 - 1
pts = (centres[torch.randint(0, 256, (P,), generator=g)] + 0.05 * torch.randn(P, dim, generator=g)).to(dev)
x = (torch.rand(Bq, H, 2 * dim, generator=g) * 2 - 1).to(dev)
cloud = ops.CloudIndex(pts)
d = torch.empty(Bq, H, device=dev)
i = torch.empty(Bq, H, device=dev, dtype=torch.int32)
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
print("dim %d Bq %d: %.2f ms  %.2f Tpairs/s  checksum %d" % (dim, Bq, ms, Bq * H * P / ms / 1e9, int(i.long().sum())))
