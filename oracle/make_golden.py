"""Generate tests/golden/*.pt by running the UNMODIFIED reference (imported from /root/reference) on CPU
with name-seeded synthetic weights (ramp_b200.testing.synthetic_state_dict) and a noise tape.

Run in the build container only:  python oracle/make_golden.py
The fixtures carry inputs + reference outputs; weights are regenerated from their seed at test time.
TEST INFRASTRUCTURE — never imported by the product."""
import contextlib
import io
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.ref_harness import NoiseTape, import_reference  # noqa: E402
from ramp_b200.testing import synthetic_cloud_2d, synthetic_cloud_3d, synthetic_state_dict  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
WEIGHT_SEED = 0


def ref_model(kind, T, compose=False, use_apf=False):
    m = import_reference()
    d3 = kind == "3d"
    with contextlib.redirect_stdout(io.StringIO()):
        unet = m.TemporalUnetInference(state_dim=6 if d3 else 4, n_support_points=64, unet_input_dim=32,
                                       dim_mults=m.UNET_DIM_MULTS[1], obstacle_3d=d3)
    cls = {"static": m.StaticGaussianDiffusionModel, "dynamic": m.DynamicGaussianDiffusionModel,
           "3d": m.GaussianDiffusionModel3d}[kind]
    model = cls(model=unet, variance_schedule="exponential", n_diffusion_steps=T, predict_epsilon=True, compose=compose,
                use_apf=use_apf, training=False)
    sd = synthetic_state_dict(model, WEIGHT_SEED)
    model.load_state_dict(sd)
    model.eval()
    for p in model.parameters():
        p.requires_grad_(False)
    return model


def hc2d():
    return {0: torch.tensor([-0.8, -0.8, 0.0, 0.0]), 63: torch.tensor([0.8, 0.8, 0.0, 0.0])}


def save(name, obj):
    os.makedirs(OUT, exist_ok=True)
    torch.save(obj, os.path.join(OUT, name + ".pt"))
    print("wrote", name, {k: (tuple(v.shape) if torch.is_tensor(v) else v) for k, v in obj.items()})


def main():
    torch.set_num_threads(8)
    m = import_reference()
    from mpd.models.diffusion_models.sample_functions import ddpm_sample_fn
    half = lambda t: 0.5  # noqa: E731

    # ---- score network rows (CFG layout of the static model: even rows conditional, odd rows zero latent)
    for kind in ("static", "3d"):
        model = ref_model(kind, 25)
        d3 = kind == "3d"
        D = 6 if d3 else 4
        g = torch.Generator("cpu").manual_seed(5)
        N = 4
        x = torch.randn(N, 64, D, generator=g)
        cloud = synthetic_cloud_3d() if d3 else synthetic_cloud_2d()
        pts = cloud[None].repeat(N, 1, 1, 1)
        t = torch.full((N,), 17)
        model.model.reset_cache()
        eps = model.model(x, t, None, obstacle_pts=pts)
        lat = model.model.cached_scene_latents.clone()
        f = model.model.forward_no_energy(x, t, obstacle_pts=pts)
        save("unet_rows_" + kind, dict(x=x, t=17, cloud=cloud, eps=eps, f=f.detach(), latent=lat[0],
                                       uncond_rows=[1, 3] if not d3 else [1], weight_seed=WEIGHT_SEED))

    # ---- static DDPM, 25 steps (BASELINE config 1 shape, B = 4)
    model = ref_model("static", 25)
    model.ddim = False
    cloud = synthetic_cloud_2d()
    tape = NoiseTape(100)
    with tape.patched():
        chain = model.run_inference(None, hc2d(), n_samples=4, horizon=64, return_chain=True,
                                    traj_normalized=torch.zeros(64, 4), obstacle_pts=cloud, sample_fn=ddpm_sample_fn,
                                    noise_std_extra_schedule_fn=half)
    save("static_ddpm25", dict(chain=chain, cloud=cloud, noise_seed=100, n_draws=tape.k, B=4, T=25,
                               weight_seed=WEIGHT_SEED))

    # ---- static DDIM, 5 of 100 steps
    model = ref_model("static", 100)
    tape = NoiseTape(200)
    with tape.patched():
        chain = model.run_inference(None, hc2d(), n_samples=4, horizon=64, return_chain=True,
                                    traj_normalized=torch.zeros(64, 4), obstacle_pts=cloud, sample_fn=ddpm_sample_fn)
    save("static_ddim5", dict(chain=chain, cloud=cloud, noise_seed=200, B=4, T=100, weight_seed=WEIGHT_SEED))

    # ---- static compose + APF, 8 of 100 DDIM steps; clouds near the clamp corners so the APF fires
    model = ref_model("static", 100, compose=True, use_apf=True)
    cloud2 = synthetic_cloud_2d(seed=3, groups=2)
    cloud2 = cloud2 * 1.3          # push obstacle boxes towards +-0.9 where clamped x0 waypoints live
    hits = []
    import mpd.models.diffusion_models.diffusion_model_static as dms
    orig_av = dms.avoidance

    def counting(traj, field, **kw):
        d, _ = field.query_static(traj[..., :2].reshape(-1, 2).cpu().numpy())
        hits.append(int((d < field.distance_threshold).sum()))
        return orig_av(traj, field, **kw)

    dms.avoidance = counting
    tape = NoiseTape(300)
    with tape.patched():
        chain = model.run_inference(None, hc2d(), n_samples=4, horizon=64, return_chain=True,
                                    traj_normalized=torch.zeros(64, 4), obstacle_pts=cloud2, sample_fn=ddpm_sample_fn)
    dms.avoidance = orig_av
    print("APF colliding waypoints per call:", hits)
    save("static_compose_apf8", dict(chain=chain, cloud=cloud2, noise_seed=300, B=4, T=100, apf_hits=hits,
                                     weight_seed=WEIGHT_SEED))

    # ---- 3-D DDPM 25 steps, B = 1 (the only batch size the reference supports)
    model = ref_model("3d", 25)
    cloud3 = synthetic_cloud_3d()
    hc = {0: torch.tensor([-0.8, -0.8, -0.8, 0.0, 0.0, 0.0]), 63: torch.tensor([0.8, 0.8, 0.8, 0.0, 0.0, 0.0])}
    tape = NoiseTape(400)
    with tape.patched():
        chain = model.run_inference(None, hc, n_samples=1, horizon=64, return_chain=True,
                                    traj_normalized=torch.zeros(64, 6), obstacle_pts=cloud3, sample_fn=ddpm_sample_fn,
                                    noise_std_extra_schedule_fn=half)
    save("3d_ddpm25", dict(chain=chain, cloud=cloud3, noise_seed=400, B=1, T=25, weight_seed=WEIGHT_SEED))

    # ---- static APF on hand-made trajectories that cross the obstacles
    from mpd.models.diffusion_models.APFhelper import ObstacleField, avoidance
    g = torch.Generator("cpu").manual_seed(9)
    cl = synthetic_cloud_2d(seed=1).reshape(-1, 2)
    s = torch.linspace(0, 1, 64)[None, :, None]
    starts = torch.tensor([[-0.8, -0.8], [-0.8, 0.8], [0.0, -0.9], [-0.9, 0.1], [0.9, 0.9]])[:, None, :]
    ends = torch.tensor([[0.8, 0.8], [0.8, -0.8], [0.0, 0.9], [0.9, -0.1], [0.95, 0.95]])[:, None, :]
    pos = starts + s * (ends - starts) + 0.01 * torch.randn(5, 64, 2, generator=g)
    traj = torch.cat([pos, 0.1 * torch.randn(5, 64, 2, generator=g)], dim=-1)
    field = ObstacleField(cl.numpy(), distance_threshold=0.07)
    d, i = field.query_static(traj[..., :2].reshape(-1, 2).numpy())
    out = avoidance(traj.clone(), field, avoidance_window=7, avoidance_strength=0.1)
    save("apf_static", dict(traj=traj, cloud=cl, out=out, dist=torch.from_numpy(d), idx=torch.from_numpy(i), thr=0.07,
                            strength=0.1, window=7))

    # ---- cost / selection
    from mpd.models.diffusion_models.cost import compute_trajectory_costs
    g = torch.Generator("cpu").manual_seed(10)
    B = 24
    starts = torch.tensor([-0.8, -0.8]) + 0.02 * torch.randn(B, 1, 2, generator=g)
    ends = torch.tensor([0.8, 0.8]) + 0.02 * torch.randn(B, 1, 2, generator=g)
    bend = 0.5 * torch.randn(B, 1, 2, generator=g) * torch.sin(np.pi * s)
    pos = starts + s * (ends - starts) + bend
    traj = torch.cat([pos, torch.diff(pos, dim=1, prepend=pos[:, :1]) * 10], dim=-1)
    cl = synthetic_cloud_2d(seed=2)
    best, bcost, costs, free, bidx = compute_trajectory_costs(traj, cl, collision_threshold=0.05)
    save("cost_select", dict(traj=traj, cloud=cl, thr=0.05, best=best, best_cost=bcost, costs=costs, free=free,
                             best_index=bidx))

    # ---- dynamic (per-trajectory) APF
    import mpd.models.diffusion_models.APFhelper_dynamic as AD
    np.random.seed(0)
    centers = np.array([[-0.4, -0.3], [0.3, 0.35], [-0.3, 0.45], [0.45, -0.4]])
    sizes = np.array([[0.3, 0.3], [0.25, 0.35], [0.3, 0.2], [0.2, 0.3]])
    pursuer = np.array([0.1, 0.05])
    fld = AD.ObstacleField(centers, sizes, lambda t, sp, rg, bi: (pursuer, 0.1), 64, distance_threshold=0.2,
                           distance_threshold_pred=0.5)
    fld.update_dynamic(0, None, replan_guide=True)
    g = torch.Generator("cpu").manual_seed(12)
    B = 6
    pos = torch.tensor([-0.8, -0.8]) + s * (torch.tensor([0.8, 0.8]) - torch.tensor([-0.8, -0.8])) \
        + 0.3 * torch.randn(B, 1, 2, generator=g) * torch.sin(np.pi * s)
    traj = torch.cat([pos, torch.zeros(B, 64, 2)], dim=-1)
    goal = traj[0, -1].clone()
    out_s, out_d = [], []
    for b in range(B):
        a = AD.avoidance(traj[b].clone(), fld, is_dynamic=False, avoidance_window=8, avoidance_strength=0.15,
                         avoidance_strength_pred=0.15)
        out_s.append(a.clone())
        c = AD.avoidance(a.clone(), fld, is_dynamic=True, avoidance_window=5, avoidance_strength=0.15,
                         avoidance_strength_pred=0.15, affected_states=64, goal_state=goal)
        out_d.append(c.clone())
    save("apf_dynamic", dict(traj=traj, static_pts=torch.from_numpy(fld.static_obstacle_points).float(),
                             disc_pts=torch.from_numpy(np.asarray(fld.dynamic_kdtree.data)).float(), goal=goal,
                             out_static=torch.stack(out_s), out_dynamic=torch.stack(out_d), thr_static=0.2, thr_pred=0.5,
                             strength=0.15, window_static=8))


if __name__ == "__main__":
    main()
