"""tests/golden/static_ddpm25_apf.pt: the UNMODIFIED reference's DDPM loop with `use_apf=True` (CFG, ddim=False): for
forward_t > 20 `p_mean_variance` pushes the posterior mean through `avoidance` (threshold 0.07, strength 0.1, window 5)
before `ddpm_sample_fn` adds the noise (diffusion_model_static.py:172-181).  Build container only
(imports /root/reference):  python oracle/make_golden_ddpm_apf.py
TEST INFRASTRUCTURE — never imported by the product."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.make_golden import hc2d, ref_model, save  # noqa: E402
from oracle.ref_harness import NoiseTape, import_reference  # noqa: E402
from ramp_b200.testing import synthetic_cloud_2d  # noqa: E402


def main():
    torch.set_num_threads(8)
    import_reference()
    from mpd.models.diffusion_models.sample_functions import ddpm_sample_fn
    import mpd.models.diffusion_models.diffusion_model_static as dms
    model = ref_model("static", 25, compose=False, use_apf=True)
    model.ddim = False
    cloud = synthetic_cloud_2d(seed=3) * 1.3          # boxes towards +-0.9, where the late (clamped) waypoints live
    hits, orig = [], dms.avoidance

    def counting(traj, field, **kw):
        d, _ = field.query_static(traj[..., :2].reshape(-1, 2).cpu().numpy())
        hits.append(int((d < field.distance_threshold).sum()))
        return orig(traj, field, **kw)

    dms.avoidance = counting
    tape = NoiseTape(500)
    with tape.patched():
        chain = model.run_inference(None, hc2d(), n_samples=4, horizon=64, return_chain=True,
                                    traj_normalized=torch.zeros(64, 4), obstacle_pts=cloud, sample_fn=ddpm_sample_fn,
                                    noise_std_extra_schedule_fn=lambda t: 0.5)
    dms.avoidance = orig
    print("APF colliding waypoints per call (forward_t = 21..24):", hits)
    assert len(hits) == 4 and sum(hits) > 0
    save("static_ddpm25_apf", dict(chain=chain, cloud=cloud, noise_seed=500, B=4, T=25, apf_hits=hits, weight_seed=0))


if __name__ == "__main__":
    main()
