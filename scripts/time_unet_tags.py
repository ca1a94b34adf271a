#!/usr/bin/env python
"""Per-launch CUDA-event times of ONE 7168-row score-network evaluation, summed by (class, tag) and printed for the tags that
match argv[1] (regex, default: the FF / LayerNorm epilogue launches at L = 64 / 32).  Honours RAMP_B200_LIB, so the same
script times diagnostic builds (csrc/gemm_tc.cu RAMP_GEMM_DIAG) on the same box."""
import collections
import csv
import re
import sys
import tempfile
import torch

sys.path.insert(0, ".")
from ramp_b200.testing import synthetic_state_dict      # noqa: E402
from ramp_b200.unet import TemporalUnetInference        # noqa: E402

pat = re.compile(sys.argv[1] if len(sys.argv) > 1 else r"L(64|32)_e[1234]")
rows = 7168
dev = torch.device("cuda", 0)
net = TemporalUnetInference(n_support_points=64, state_dim=4, max_rows=rows, max_scenes=7)
net.load_state_dict(synthetic_state_dict(net, 0))
net = net.to(dev)
net.prepare(horizon=64, max_rows=rows, max_scenes=7)
g = torch.Generator("cpu").manual_seed(0)
net.set_scene_latents((torch.randn(7, net.context_dim, generator=g) * 0.5).to(dev))
x = torch.randn(rows, 64, 4, generator=g).to(dev)
r2s = (torch.arange(rows) % 7).to(torch.int32).to(dev)
for _ in range(3):
    net.energy_grad_rows(x, 12, r2s)
torch.cuda.synchronize()
net.profile(True)
for _ in range(3):
    net.energy_grad_rows(x, 12, r2s)
torch.cuda.synchronize()
path = tempfile.mktemp(suffix=".csv")
net.profile_dump(path)
net.profile(False)
agg = collections.OrderedDict()
for r in csv.DictReader(open(path)):
    a = agg.setdefault((r["class"], r["tag"]), [0, 0.0])
    a[0] += 1
    a[1] += float(r["ms"])
tot = collections.Counter()
for (c, t), (n, ms) in agg.items():
    tot[c] += ms / 3
    if pat.search(t):
        print("%-6s %-34s n %3d  %8.1f us/launch" % (c, t, n // 3, ms / n * 1e3))
print("per call:", {k: round(v, 2) for k, v in tot.items()}, "sum %.2f ms" % sum(tot.values()))
