"""DynamicGaussianDiffusionModel (CUDA path) vs the reference's own outputs (tests/golden/dynamic_tick.pt) and vs the
CPU oracle: stage I, teacher-forced replan ticks (q_sample, pins, 5 DDIM steps, static + pursuer APF, smoothing),
selection, and the full run_inference protocol on a synthetic pursuit environment."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")


def _model(cuda, max_rows=64):
    from ramp_b200.dynamic import DynamicGaussianDiffusionModel
    from ramp_b200.testing import synthetic_state_dict
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=64, state_dim=4, max_rows=max_rows)
    m = DynamicGaussianDiffusionModel(model=net, n_diffusion_steps=100, predict_epsilon=True)
    m.load_state_dict(synthetic_state_dict(m, 0))
    return m.to(cuda).eval()


def _hc(dev):
    return {0: torch.tensor([-0.8, -0.8, 0.0, 0.0], device=dev), 63: torch.tensor([0.8, 0.8, 0.0, 0.0], device=dev)}


def test_stage1_and_ticks_vs_reference(cuda):
    from ramp_b200 import ops
    from ramp_b200.apf_dynamic import ObstacleField
    from ramp_b200.cost import compute_trajectory_costs
    from ramp_b200.testing import tape
    g = torch.load(os.path.join(G, "dynamic_tick.pt"), weights_only=False)
    m = _model(cuda)
    B = g["B"]
    cloud = g["cloud"].to(cuda)
    draws = tape(g["noise_seed"], [(B, 64, 4)] * 8)
    # stage I: 10 DDIM steps, no APF
    hcB = m._repeat_hard_conds(_hc(cuda), B)
    pins = ops.Pins.from_hard_conds(hcB, B, 64, 4, cuda)
    x = draws[0].to(cuda).contiguous()
    ops.inpaint(x, pins)
    for i in m.ddim_set_timesteps(10):
        m._ddim_step(x, int(i), int(i) - 10, cloud, pins=pins, out=x)
    err = (x.cpu() - g["stage1_batch"]).abs().max().item()
    med = (x.cpu() - g["stage1_batch"]).abs().median().item()
    print("stage I (10 free-running DDIM steps) max err %.2e median %.2e" % (err, med))
    assert err < 1e-3 and med < 1e-4       # 10 free-running steps within north_star's 1e-3
    best, _, _, free, _ = compute_trajectory_costs(g["stage1_batch"].to(cuda), cloud, collision_threshold=0.02)
    assert (best.cpu() - g["stage1_best"]).abs().max() < 1e-6
    # stage II, teacher-forced on the reference's selected plans
    np.random.seed(0)
    field = ObstacleField(np.zeros((1, 2)), np.ones((1, 2)), None, 8, distance_threshold=0.2, distance_threshold_pred=0.5,
                          device=cuda)
    field.static_obstacle_points = g["static_pts"].numpy()
    field.static_cloud = ops.CloudIndex(g["static_pts"].to(cuda))
    xb = g["stage1_best"].clone()
    history = [xb[0].clone().unsqueeze(0).to(cuda)]
    for k in range(g["n_ticks"]):
        t = g["tick%d" % k]
        field.set_dynamic_points(t["disc"], t["center"].numpy())
        batch = m._tick_batch(xb.to(cuda).contiguous(), history, k, k, B, cloud, field, noise=draws[t["draw_index"]].to(cuda))
        err = (batch.cpu() - t["batch"]).abs().max().item()
        print("tick %d batch max err %.2e" % (k, err))
        assert err < 1e-3, (k, err)
        best, _, _, free, bidx = compute_trajectory_costs(t["batch"].to(cuda), t["cost_cloud"].to(cuda), collision_threshold=0.05)
        assert torch.equal(free.cpu(), t["free"]) and (best.cpu() - t["best"]).abs().max() < 1e-6
        xb = t["best"].clone()
        xb[0, 2:] = 0.0
        history.append(xb[k + 1].clone().unsqueeze(0).to(cuda))


def test_row_scene_quirk_matches_oracle(cuda):
    """Even and odd batches: the stacked-CFG row parity of the reference (UnetInference.py:192-195)."""
    from oracle import ramp_oracle as O
    from ramp_b200.testing import pursuit_cloud_2d, synthetic_state_dict, SyntheticPursuitDataset
    m = _model(cuda)
    sd = synthetic_state_dict(m, 0)
    cloud = pursuit_cloud_2d(SyntheticPursuitDataset())
    lat = O.scene_encoder_2d(sd, cloud[None])[0]
    for B in (4, 3):
        gen = torch.Generator("cpu").manual_seed(B)
        x = torch.randn(B, 64, 4, generator=gen)
        eps = torch.empty(B, 64, 4, device=cuda)
        m._x0_step(x.to(cuda), 40, cloud.to(cuda), eps_out=eps)
        ref = O.eps_dynamic(sd, x, 40, lat)
        err = (eps.cpu() - ref).abs().max().item()
        print("B=%d eps err %.2e" % (B, err))
        assert err < 5e-4


def test_run_inference_protocol(cuda):
    """Full two-stage loop on the synthetic pursuit environment: shapes, pins, context side effects."""
    from ramp_b200.sample_functions import ddpm_sample_fn
    from ramp_b200.testing import pursuit_cloud_2d, SyntheticPursuitDataset
    m = _model(cuda)
    m.max_iteration = 3
    ds = SyntheticPursuitDataset(device=cuda)
    cloud = pursuit_cloud_2d(ds).to(cuda)
    ctx = {'dataset': ds}
    np.random.seed(0)
    torch.manual_seed(0)
    chain, chain_obs, chain_start = m.run_inference(ctx, _hc(cuda), n_samples=5, horizon=64, return_chain=True,
                                                    traj_normalized=torch.zeros(64, 4, device=cuda), obstacle_pts=cloud,
                                                    sample_fn=ddpm_sample_fn)
    assert chain.shape == (4, 1, 64, 4) and len(chain_obs) == 3 and len(chain_start) == 4
    assert torch.isfinite(chain).all()
    # executed history is pinned: tick j's plan starts with the states executed so far
    for j in range(1, 3):
        assert torch.allclose(chain[j + 1, 0, :j, :2], chain[j, 0, :j, :2], atol=1e-6)
    assert torch.allclose(chain[-1, 0, -1], torch.tensor([0.8, 0.8, 0.0, 0.0], device=cuda))
    assert 'obstacle_field' not in ctx          # run_inference works on a shallow copy of the context (:650)


def test_replan_graph_streams_the_conditioning_cloud(cuda):
    """ReplanGraph(stream_conditioning=True): after a replay the engine is conditioned on whatever the static `cloud`
    buffer held at that replay (encoder kernels + scene constants are graph nodes), bit-equal to an eager re-encode."""
    import numpy as np
    from ramp_b200.apf_dynamic import generate_box_points, generate_sphere_points
    from ramp_b200.graphs import ReplanGraph
    from ramp_b200.testing import SyntheticPursuitDataset, pursuit_cloud_2d
    m = _model(cuda)
    ds = SyntheticPursuitDataset()
    cloud_a = pursuit_cloud_2d(ds).to(cuda)
    g0 = torch.Generator("cpu").manual_seed(2)
    cloud_b = (cloud_a.cpu() + 0.05 * torch.randn(cloud_a.shape, generator=g0)).to(cuda)
    np.random.seed(0)
    boxes = ds.env.obj_fixed_list[0].fields[0]
    static_pts = torch.from_numpy(np.vstack([generate_box_points(c, s, 64) for c, s in
                                             zip(boxes.centers.numpy(), boxes.sizes.numpy())])).float()
    disc = torch.from_numpy(generate_sphere_points(np.array([-0.35, -0.95]), 0.1, 64)).float()
    H, k, B = 64, 2, 8
    s = torch.linspace(0, 1, H)[:, None]
    plan = torch.cat([-0.8 + 1.6 * s, -0.8 + 1.6 * s, torch.full((H, 2), 0.2)], dim=1).to(cuda)
    g = ReplanGraph(m, B, H, k, cloud_a, static_pts, disc, static_pts, stream_conditioning=True)
    g.x_clean.copy_(plan)
    g.history.copy_(plan[:k + 1])
    g.disc_center.copy_(torch.tensor([-0.35, -0.95]))
    x = torch.randn(2 * B, H, 4, generator=torch.Generator("cpu").manual_seed(1)).to(cuda)
    r2s = m._row_scenes(B, 1, cuda)
    seen = []
    for cl in (cloud_a, cloud_b):
        g.cloud.copy_(cl)
        g.replay()
        torch.cuda.synchronize()
        assert torch.isfinite(g.batch).all()
        seen.append(m.model.energy_grad_rows(x, 10, r2s).clone())            # engine state left by the replay
    assert not torch.equal(seen[0], seen[1])
    for cl, got in zip((cloud_a, cloud_b), seen):
        m._scene_sig = None
        m._set_scenes(m._groups_and_weights(cl, False)[0])                  # eager re-encode
        assert torch.equal(m.model.energy_grad_rows(x, 10, r2s), got)
