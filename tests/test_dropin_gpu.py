"""Reference-shaped drop-in modules (cost.py, APFhelper.py, APFhelper_dynamic.py names) on the CUDA path vs the
reference's own outputs (tests/golden) — the calls read like the reference's call sites."""
import os

import numpy as np
import pytest
import torch

gpu = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")


def load(name):
    return torch.load(os.path.join(G, name + ".pt"), weights_only=False)


@gpu
def test_compute_trajectory_costs_matches_reference(cuda):
    from ramp_b200.cost import (compute_collision_with_pointcloud, compute_path_length, compute_smoothness,
                                compute_trajectory_costs)
    g = load("cost_select")
    trajs, cloud = g["traj"].to(cuda), g["cloud"].to(cuda)
    best, bcost, costs, free, bidx = compute_trajectory_costs(trajs, cloud, collision_threshold=g["thr"])
    assert torch.equal(free.cpu(), g["free"])
    assert int(bidx) == int(g["best_index"])
    assert (best.cpu() - g["best"]).abs().max() == 0
    assert abs(float(bcost) - float(g["best_cost"])) < 1e-5
    assert costs.shape == g["costs"].shape and (costs.cpu() - g["costs"]).abs().max() < 1e-5
    assert torch.equal(compute_collision_with_pointcloud(trajs, cloud, g["thr"]).cpu(), ~g["free"])
    ref_len = torch.linalg.norm(torch.diff(g["traj"][:, :, :2], dim=-2), dim=-1).sum(-1)
    ref_sm = torch.linalg.norm(torch.diff(g["traj"][:, :, 2:], dim=-2), dim=-1).sum(-1)
    assert (compute_path_length(trajs).cpu() - ref_len).abs().max() < 1e-4
    assert (compute_smoothness(trajs).cpu() - ref_sm).abs().max() < 1e-4


@gpu
def test_compute_trajectory_costs_none_protocol(cuda, capsys):
    """cost.py:59-61: every trajectory collides -> (None, None, None, mask, None) and the reference's message."""
    from ramp_b200.cost import compute_trajectory_costs
    g = load("cost_select")
    out = compute_trajectory_costs(g["traj"].to(cuda), g["cloud"].to(cuda), collision_threshold=10.0)
    assert out[0] is None and out[1] is None and out[2] is None and out[4] is None
    assert not bool(out[3].any())
    assert "No collision-free" in capsys.readouterr().out


@gpu
def test_static_obstacle_field_and_avoidance(cuda):
    from ramp_b200.apf import ObstacleField, avoidance
    g = load("apf_static")
    field = ObstacleField(g["cloud"].numpy(), distance_threshold=g["thr"], device=cuda)
    d, i = field.query_static(g["traj"][..., :2].reshape(-1, 2).to(cuda))
    assert np.array_equal(i.cpu().numpy(), g["idx"].numpy())
    fin = torch.isfinite(g["dist"])
    assert torch.equal(torch.isfinite(d.cpu()), fin)
    assert (d.cpu()[fin].double() - g["dist"][fin]).abs().max() < 1e-6
    x = g["traj"].to(cuda)
    out = avoidance(x, field, avoidance_window=g["window"], avoidance_strength=g["strength"])
    assert out.data_ptr() != x.data_ptr()                       # the reference returns a new tensor
    assert (out.cpu() - g["out"]).abs().max() < 1e-5
    assert torch.equal(out[..., 2:].cpu(), g["traj"][..., 2:])  # velocities untouched


@gpu
def test_dynamic_obstacle_field_and_avoidance(cuda):
    from ramp_b200 import ops
    from ramp_b200.apf_dynamic import ObstacleField, avoidance
    g = load("apf_dynamic")
    np.random.seed(0)
    field = ObstacleField(np.array([[0.0, 0.0]]), np.array([[0.1, 0.1]]), lambda t, sp, rg, bi: (np.zeros(2), 0.1), 8,
                          distance_threshold=g["thr_static"], distance_threshold_pred=g["thr_pred"], device=cuda)
    field.static_cloud = ops.CloudIndex(g["static_pts"].to(cuda))
    field.set_dynamic_points(g["disc_pts"], np.zeros(2))
    B = g["traj"].shape[0]
    for b in range(B):
        t = g["traj"][b].to(cuda).contiguous().clone()
        avoidance(t, field, is_dynamic=False, avoidance_window=g["window_static"], avoidance_strength=g["strength"],
                  avoidance_strength_pred=g["strength"])
        assert (t.cpu() - g["out_static"][b]).abs().max() < 1e-5
        avoidance(t, field, is_dynamic=True, avoidance_window=5, avoidance_strength=g["strength"],
                  avoidance_strength_pred=g["strength"], affected_states=64, goal_state=g["goal"].to(cuda))
        assert (t.cpu() - g["out_dynamic"][b]).abs().max() < 1e-5


def test_point_generators_follow_reference_rng_order():
    """generate_box_points / generate_sphere_points draw from numpy's global RNG in the reference's order
    (APFhelper_dynamic.py:18-68): same seed -> the clouds committed in the golden file."""
    from ramp_b200.apf_dynamic import generate_box_points, generate_sphere_points
    g = load("apf_dynamic")
    np.random.seed(0)
    centers = np.array([[-0.4, -0.3], [0.3, 0.35], [-0.3, 0.45], [0.45, -0.4]])
    sizes = np.array([[0.3, 0.3], [0.25, 0.35], [0.3, 0.2], [0.2, 0.3]])
    static = np.vstack([generate_box_points(c, s, 64) for c, s in zip(centers, sizes)])
    disc = generate_sphere_points(np.array([0.1, 0.05]), 0.1, 64)
    assert np.abs(static - g["static_pts"].numpy()).max() < 1e-6
    assert np.abs(disc - g["disc_pts"].numpy()).max() < 1e-6
