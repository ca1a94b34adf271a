#!/usr/bin/env python
"""Profiling driver: the score network at the headline chunk (7168 rows = 1024 trajectories x 7 score rows), two
energy-gradient calls (the first is warm-up).  Run under ncu with --kernel-name / --launch-skip to pick a launch;
gemm launch indices inside one call come from `bench.py --dump-launches` (260 gemm launches per call)."""
import sys
import torch

sys.path.insert(0, ".")
from ramp_b200.testing import synthetic_state_dict      # noqa: E402
from ramp_b200.unet import TemporalUnetInference        # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 7168
calls = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dev = torch.device("cuda", 0)
net = TemporalUnetInference(n_support_points=64, state_dim=4, max_rows=rows, max_scenes=7)
net.load_state_dict(synthetic_state_dict(net, 0))
net = net.to(dev)
net.prepare(horizon=64, max_rows=rows, max_scenes=7)
g = torch.Generator("cpu").manual_seed(0)
net.set_scene_latents((torch.randn(7, net.context_dim, generator=g) * 0.5).to(dev))
x = torch.randn(rows, 64, 4, generator=g).to(dev)
r2s = (torch.arange(rows) % 7).to(torch.int32).to(dev)
for _ in range(calls):
    eps = net.energy_grad_rows(x, 12, r2s)
torch.cuda.synchronize()
print("ok", float(eps.abs().max()))
