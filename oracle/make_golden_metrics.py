"""Generate tests/golden/metrics.pt with the UNMODIFIED reference metrics (scripts/inference/core/metrics.py, loaded
from /root/reference by path; it only needs torch + numpy) on seeded synthetic batches.  Build container only."""
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference/scripts/inference/core/metrics.py"


def smooth_batch(n, H, seed):
    """Random-walk trajectories from (-0.8, -0.8) towards (0.8, 0.8) with velocities (the shape of a sampled batch)."""
    g = torch.Generator("cpu").manual_seed(seed)
    t = torch.linspace(0, 1, H)[None, :, None]
    pos = -0.8 + 1.6 * t + 0.25 * torch.cumsum(torch.randn(n, H, 2, generator=g), dim=1) / H ** 0.5 * (t * (1 - t) * 4)
    vel = torch.diff(pos, dim=1, prepend=pos[:, :1]) * 10
    return torch.cat([pos, vel], dim=-1).contiguous()


def main():
    spec = importlib.util.spec_from_file_location("ref_metrics", REF)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    M, DM = ref.Metrics(), ref.DynamicMetrics()
    g = torch.Generator("cpu").manual_seed(11)
    centers = torch.rand(6, 2, generator=g) * 1.4 - 0.7
    sizes2 = torch.rand(6, 2, generator=g) * 0.2 + 0.1
    sizes1 = torch.full((6,), 0.26)
    cases = []
    for n, seed in ((1, 1), (2, 2), (37, 3), (300, 4)):
        x = smooth_batch(n, 64, seed)
        for sizes in (sizes1, sizes2):
            ci = M.compute_collision_intensity(x, centers, sizes)
            m = M.trajectory_success_and_metrics(x, ci, threshold=0.01)
            m = {k: v for k, v in m.items() if k != "free_trajectories"}
            cases.append(dict(x=x, centers=centers, sizes=sizes, ci=ci, metrics=m, path_length=M.compute_path_length(x),
                              smoothness=M.compute_smoothness(x),
                              variance_all=M.compute_variance_waypoints(x) if n > 1 else None))
    rng = np.random.default_rng(5)
    ev = [np.array([-0.8, -0.8]) + 0.05 * k + 0.01 * rng.standard_normal(2) for k in range(30)]
    pu = [np.array([-0.35, -0.95]) + 0.03 * k * np.array([0.6, 0.8]) for k in range(28)]
    episodes = []
    for static_collision, radius, goal_thr in ((False, 0.05, 0.2), (True, 0.05, 0.2), (False, 0.6, 0.2), (False, 0.05, 5.0)):
        out = DM.calculate_single_episode_metrics(ev, pu, torch.tensor([-0.8, -0.8]), torch.tensor([0.8, 0.8]), goal_thr,
                                                  static_collision, radius)
        episodes.append(dict(args=(static_collision, radius, goal_thr), out=out))
    path = os.path.join(ROOT, "tests", "golden", "metrics.pt")
    torch.save(dict(cases=cases, evader=ev, pursuer=pu, episodes=episodes), path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
