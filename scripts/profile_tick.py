#!/usr/bin/env python
"""Where a replan tick's time goes: one EAGER stage-II tick of the dynamic planner at B = 1024 (the
shape of bench.py's `replan` section: 5 DDIM steps, 4096-point clouds) with CUDA events around every launch of the score
network (ramp_unet_profile), aggregated by sequence length L and kernel class; the non-network tail (q_sample, APF,
smoothing, cost, selection) is timed as one block.     python scripts/profile_tick.py > profiles/rN_tick_levels.txt"""
import collections
import csv
import os
import re
import sys
import tempfile

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ramp_b200 import ops  # noqa: E402
from ramp_b200.apf_dynamic import generate_box_points, generate_sphere_points  # noqa: E402
from ramp_b200.dynamic import DynamicGaussianDiffusionModel  # noqa: E402
from ramp_b200.testing import SyntheticPursuitDataset, pursuit_cloud_2d, synthetic_state_dict  # noqa: E402
from ramp_b200.unet import TemporalUnetInference  # noqa: E402

B, H, D, k = int(sys.argv[1]) if len(sys.argv) > 1 else 1024, 64, 4, 3
dev = torch.device("cuda:0")
net = TemporalUnetInference(n_support_points=H, state_dim=D, max_rows=2 * B, max_scenes=2)
m = DynamicGaussianDiffusionModel(model=net, n_diffusion_steps=100, predict_epsilon=True)
m.load_state_dict(synthetic_state_dict(m, 0))
m = m.to(dev).eval()
ds = SyntheticPursuitDataset()
cloud = pursuit_cloud_2d(ds).to(dev)
np.random.seed(0)
boxes = ds.env.obj_fixed_list[0].fields[0]
static_pts = torch.from_numpy(np.vstack([generate_box_points(c, s, 960) for c, s in zip(boxes.centers.numpy(), boxes.sizes.numpy())])).float()
disc = torch.from_numpy(generate_sphere_points(np.array([-0.35, -0.95]), 0.1, 256)).float()


class Field:
    pass


f = Field()
f.static_cloud = ops.CloudIndex(static_pts.to(dev))
f.dynamic_cloud = ops.CloudIndex(disc.to(dev))
f.dynamic_center_dev = torch.tensor([-0.35, -0.95], device=dev)
f.distance_threshold, f.distance_threshold_pred = m.distance_threshold_static, m.distance_threshold_pred_apf
s = torch.linspace(0, 1, H)[:, None]
plan = torch.cat([-0.8 + 1.6 * s, -0.8 + 1.6 * s, torch.full((H, 2), 0.2)], dim=1).to(dev)
history = [plan[j].clone().unsqueeze(0) for j in range(k + 1)]
cost_cloud = ops.CloudIndex(torch.cat([static_pts, disc], 0).to(dev))


def tick():
    batch = m._tick_batch(plan, history, k, k, B, cloud, f)
    return ops.select_best(ops.cost_terms(batch, cost_cloud, m.collision_threshold_low))


for _ in range(3):
    tick()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for _ in range(5):
    tick()
e1.record()
torch.cuda.synchronize()
eager_ms = e0.elapsed_time(e1) / 5
m.model.profile(True)
tick()
path = os.path.join(tempfile.gettempdir(), "tick_launches.csv")
m.model.profile_dump(path)
ms, cnt = m.model.profile(False, fetch=True)
agg = collections.defaultdict(lambda: [0, 0.0])
for r in csv.DictReader(open(path)):
    mL = re.search(r"_L(\d+)", r["tag"])
    L = int(mL.group(1)) if mL else 0
    kind = r["class"]
    if kind == "gemm":
        e = re.search(r"_e(\d)", r["tag"])
        kind = "gemm_e%s" % (e.group(1) if e else "?")
    agg[(L, kind)][0] += 1
    agg[(L, kind)][1] += float(r["ms"])
net_ms = sum(v[1] for v in agg.values())
print("# one eager replan tick, B = %d (network rows per step %d), 5 DDIM steps: %.2f ms per tick (eager, device time)" % (B, B, eager_ms))
print("# score-network launches: %d, %.2f ms (sum of per-launch CUDA-event times); everything else (q_sample, pins, APF, smoothing, cost, selection, launch gaps): %.2f ms"
      % (sum(v[0] for v in agg.values()), net_ms, eager_ms - net_ms))
print("# gemm_e0 plain, e1 GEGLU fwd, e2 FF2^T + GEGLU bwd, e3 LayerNorm fwd, e4 LayerNorm bwd epilogues")
print("%-4s %-12s %8s %10s %9s %8s" % ("L", "class", "launches", "total_ms", "mean_us", "share"))
for (L, kind), (n, t) in sorted(agg.items(), key=lambda kv: (-kv[0][0], kv[0][1])):
    print("%-4d %-12s %8d %10.3f %9.1f %7.1f%%" % (L, kind, n, t, 1e3 * t / n, 100 * t / net_ms))
byL = collections.defaultdict(float)
for (L, _), (_, t) in agg.items():
    byL[L] += t
print("# per level:", "  ".join("L=%d %.2f ms (%.0f%%)" % (L, t, 100 * t / net_ms) for L, t in sorted(byL.items(), reverse=True)))
