"""Generate tests/golden/verdicts.pt: north_star's second parity criterion — "identical collision-free / success verdicts on
a fixed seed set".  The UNMODIFIED reference (StaticGaussianDiffusionModel, 25 DDPM steps, BASELINE config-1 shape) samples
B trajectories for each seed of a fixed set; its own Metrics.compute_collision_intensity / trajectory_success_and_metrics
(scripts/inference/core/metrics.py:48-126) and compute_collision_with_pointcloud / compute_trajectory_costs
(mpd/models/diffusion_models/cost.py:25-88) give the verdicts.  Build container only; test infrastructure."""
import importlib.util
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

from make_golden import hc2d, ref_model  # noqa: E402
from oracle.ref_harness import NoiseTape, import_reference  # noqa: E402
from ramp_b200.testing import synthetic_cloud_2d  # noqa: E402

SEEDS = (1000, 1100, 1200, 1300, 1400, 1500)
B = 6


def main():
    torch.set_num_threads(8)
    import_reference()
    from mpd.models.diffusion_models.sample_functions import ddpm_sample_fn
    from mpd.models.diffusion_models.cost import compute_collision_with_pointcloud, compute_trajectory_costs
    spec = importlib.util.spec_from_file_location("ref_metrics", "/root/reference/scripts/inference/core/metrics.py")
    RM = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(RM)
    M = RM.Metrics()
    model = ref_model("static", 25)
    model.ddim = False
    cases = []
    for i, seed in enumerate(SEEDS):
        cloud = synthetic_cloud_2d(seed=i)
        g = torch.Generator("cpu").manual_seed(1000 + i)                      # the box centres synthetic_cloud_2d drew first
        centres = ((torch.rand(6, 1, 2, generator=g) * 2 - 1) * 0.7).reshape(6, 2)
        sizes = torch.full((6,), 0.26)
        model.model.reset_cache()
        tape = NoiseTape(seed)
        with tape.patched():
            chain = model.run_inference(None, hc2d(), n_samples=B, horizon=64, return_chain=True,
                                        traj_normalized=torch.zeros(64, 4), obstacle_pts=cloud, sample_fn=ddpm_sample_fn,
                                        noise_std_extra_schedule_fn=lambda t: 0.5)
        final = chain[-1]
        ci = M.compute_collision_intensity(final, centres, sizes)
        met = M.trajectory_success_and_metrics(final, ci, threshold=0.01)
        cloud_hit = {thr: compute_collision_with_pointcloud(final, cloud, collision_threshold=thr) for thr in (0.02, 0.05)}
        best = compute_trajectory_costs(final, cloud, collision_threshold=0.02)
        cases.append(dict(seed=seed, cloud_seed=i, cloud=cloud, centres=centres, sizes=sizes, final=final.clone(), x1=chain[1].clone(),
                          collision_intensity=ci, free=(ci <= 0.01), success=met["success"],
                          n_free=met["n_free_trajectories"], cloud_hit=cloud_hit,
                          best_index=None if best[0] is None else int(best[4]),
                          survivors=best[3]))
        print(seed, "free", (ci <= 0.01).tolist(), "hit.02", cloud_hit[0.02].tolist(), "best", cases[-1]["best_index"])
    path = os.path.join(ROOT, "tests", "golden", "verdicts.pt")
    torch.save(dict(cases=cases, B=B, T=25, weight_seed=0), path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
