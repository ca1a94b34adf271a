"""SURVEY §8f f4 — the pursuer's motion and the obstacle samplers.  CPU: the oracle restatement (numpy, draws as inputs)
and the host drop-ins against what the UNMODIFIED reference produced (tests/golden/pursuit.pt, oracle/
make_golden_pursuit.py).  GPU: the kernels behind ramp_pursuer_tick / ramp_box_points against the same vectors and
against the oracle at other sizes, and the device pursuer as the first node of a graph-replayed replan tick.

Tolerances: the oracle is bit-equal to the reference (same numpy operations).  The kernels round every operation the same
way but use CUDA's double-precision sin / cos (≤ 2 ulp, glibc ≤ 1 ulp), so stored float32 points may differ by one
float32 ulp: 1.2e-7 absolute for |coordinates| ≤ 1.5."""
import os

import numpy as np
import pytest
import torch

G = os.path.join(os.path.dirname(__file__), "golden", "pursuit.pt")
ULP = 1.2e-7


@pytest.fixture(scope="module")
def gold():
    return torch.load(G, weights_only=False)


# ---------------------------------------------------------------------------------------------- CPU: oracle + host
def test_oracle_samplers_bit_equal_to_reference(gold):
    from oracle import ramp_oracle as O
    for case in gold["discs"]:
        for k, c in enumerate(case["centres"]):
            out = O.disc_points(c, case["radius"], case["num_points"], case["tape"][k])
            assert np.array_equal(out, case["points"][k]), ("disc", case["seed"], k)
    for case in gold["boxes"]:
        n = case["num_points"]
        out = np.vstack([O.box_points(case["centres"][b], case["sizes"][b], n, int(case["n_edge"][b]), case["u_edge"][b],
                                      case["u_in"][b]) for b in range(4)])
        assert np.array_equal(out, case["points"]), ("box", case["seed"])


def test_oracle_pursuer_step_bit_equal_to_reference_field(gold):
    from oracle import ramp_oracle as O
    for case in gold["tracks"]:
        _, ps, rs = case["args"]
        c = case["track"][0].numpy()
        for t in range(60):
            c = O.pursuer_step(t, c, case["robot"][t].numpy(), case["velocity"], ps, rs)
            assert np.array_equal(c, case["track"][t + 1].numpy()), (case["args"], t)


def test_host_dropins_reproduce_reference_under_the_same_seed(gold):
    """apf_dynamic.generate_* draw from np.random in the reference's order; datasets.PursuerField moves like the reference's."""
    from ramp_b200.apf_dynamic import generate_box_points, generate_sphere_points
    from ramp_b200.datasets import DynamicsGenerator, PursuerField
    for case in gold["discs"]:
        np.random.seed(case["seed"])
        for k, c in enumerate(case["centres"]):
            assert np.array_equal(generate_sphere_points(c, case["radius"], case["num_points"]), case["points"][k])
    for case in gold["boxes"]:
        np.random.seed(case["seed"])
        out = np.vstack([generate_box_points(c, s, case["num_points"]) for c, s in zip(case["centres"], case["sizes"])])
        assert np.array_equal(out, case["points"])
    for case in gold["tracks"]:
        fn, vel = DynamicsGenerator.create_pursuit_dynamics(*case["args"])
        f = PursuerField(np.array([[-0.35, -0.95]]), np.array([0.1]), dynamics_fn=fn, velocity=vel,
                         tensor_args=dict(device="cpu", dtype=torch.float32))
        for t in range(60):
            f.update_centers(t, case["robot"][t:t + 1])
            assert torch.equal(f.centers, case["track"][t + 1:t + 2]), (case["args"], t)


def test_tapes_follow_the_reference_draw_order(gold):
    from ramp_b200.pursuit import draw_box_tape, draw_disc_tape
    case = gold["discs"][0]
    np.random.seed(case["seed"])
    assert np.array_equal(draw_disc_tape(len(case["centres"]), case["num_points"]), case["tape"])
    case = gold["boxes"][0]
    np.random.seed(case["seed"])
    n_edge, u_edge, u_in = draw_box_tape(4, case["num_points"])
    assert np.array_equal(n_edge, case["n_edge"]) and np.array_equal(u_edge, case["u_edge"]) and np.array_equal(u_in, case["u_in"])


def test_device_pursuer_has_no_cpu_path():
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from ramp_b200._lib import RampError
    from ramp_b200.pursuit import DevicePursuer, box_points
    with pytest.raises(RampError):
        DevicePursuer([0.0, 0.0], device="cpu")
    with pytest.raises(RampError):
        box_points(np.zeros((1, 2), np.float32), np.ones((1, 2), np.float32), 8, device="cpu")


# ---------------------------------------------------------------------------------------------- GPU: the kernels
@pytest.mark.gpu
def test_disc_kernel_vs_reference_and_oracle(cuda, gold):
    from oracle import ramp_oracle as O
    from ramp_b200.pursuit import DevicePursuer
    for case in gold["discs"]:
        p = DevicePursuer(case["centres"][0], radius=case["radius"], num_points=case["num_points"], tape=case["tape"], device=cuda)
        robot = torch.zeros(2, device=cuda)
        for k, c in enumerate(case["centres"]):
            p.center.copy_(torch.from_numpy(c))
            pts = p.tick(robot, step=False).cpu().numpy()             # points only; the counter walks the tape
            ref = case["points"][k]
            assert np.abs(pts - ref.astype(np.float32)).max() <= ULP, (case["seed"], k)
            assert np.abs(pts - O.disc_points(c, case["radius"], case["num_points"], case["tape"][k])).max() <= 2 * ULP
        assert int(p.counter) == len(case["centres"])
    # a size the reference scripts never use: 4096 points, tape wraps around
    rs = np.random.RandomState(9)
    tape = np.concatenate((rs.uniform(0, 1, (3, 410)), rs.uniform(0, 2 * np.pi, (3, 410))), axis=1)
    p = DevicePursuer([0.3, -0.2], radius=0.3, num_points=4096, tape=tape, device=cuda)
    for k in range(5):
        pts = p.tick(torch.zeros(2, device=cuda), step=False).cpu().numpy()
        want = O.disc_points(np.float32([0.3, -0.2]), 0.3, 4096, tape[k % 3])
        assert np.abs(pts - want).max() <= 2 * ULP


@pytest.mark.gpu
def test_box_kernel_vs_reference(cuda, gold):
    from ramp_b200.pursuit import box_points
    for case in gold["boxes"]:
        out = box_points(case["centres"], case["sizes"], case["num_points"], tape=(case["n_edge"], case["u_edge"], case["u_in"]),
                         device=cuda).cpu().numpy()
        # no transcendental on this path: every operation is correctly rounded -> equal after the float32 store
        assert np.array_equal(out, case["points"].astype(np.float32)), case["seed"]
    np.random.seed(3)
    assert box_points(gold["boxes"][0]["centres"], gold["boxes"][0]["sizes"], 33, device=cuda).shape == (4 * 33, 2)


@pytest.mark.gpu
def test_pursuer_law_kernel_vs_reference_field(cuda, gold):
    from ramp_b200.pursuit import DevicePursuer
    for case in gold["tracks"]:
        vmax, ps, rs = case["args"]
        robot = case["robot"].to(cuda)
        # (a) explicit-time protocol of the reference field
        p = DevicePursuer(case["track"][0].numpy(), velocity_max=vmax, pursuit_strength=ps, random_strength=rs, device=cuda)
        worst = 0.0
        for t in range(60):
            p.update_centers(t, robot[t:t + 1])
            worst = max(worst, (p.centers.cpu() - case["track"][t + 1:t + 2]).abs().max().item())
            p.center.copy_(case["track"][t + 1].to(cuda))            # teacher-forced: compare one step at a time
        assert worst <= ULP, (case["args"], worst)
        # (b) free-running with the device counter, centre + points written to caller buffers
        p = DevicePursuer(case["track"][0].numpy(), velocity_max=vmax, pursuit_strength=ps, random_strength=rs, device=cuda)
        pts, c_out = torch.empty(64, 2, device=cuda), torch.empty(2, device=cuda)
        for t in range(60):
            p.tick(robot[t], points=pts, center_out=c_out)
        assert (c_out.cpu() - case["track"][60]).abs().max().item() <= 60 * ULP
        assert torch.equal(c_out, p.center)
        rim = (pts[:57] - c_out).norm(dim=1)
        assert (rim - 0.1).abs().max().item() < 1e-6 and ((pts[57:] - c_out).norm(dim=1) <= 0.1 + 1e-6).all()
    # pursuer on top of the robot: the reference skips the normalisation (no division by zero)
    p = DevicePursuer([0.2, 0.2], device=cuda)
    p.update_centers(3, torch.tensor([[0.2, 0.2]]))
    assert torch.isfinite(p.center).all()


def _replay_seeded(g, seed):
    torch.manual_seed(seed)
    torch.cuda.manual_seed(seed)
    g.replay()
    torch.cuda.synchronize()
    return g.batch.clone()


@pytest.mark.gpu
def test_replan_graph_with_device_pursuer_matches_host_fed_graph(cuda):
    """The same tick graph, once with the pursuer cloud computed by the host drop-ins and uploaded, once with the
    DevicePursuer node inside the graph: selected plans agree, and the pursuer moves on every replay without host work."""
    from oracle import ramp_oracle as O
    from ramp_b200.dynamic import DynamicGaussianDiffusionModel
    from ramp_b200.graphs import ReplanGraph
    from ramp_b200.pursuit import DevicePursuer, draw_disc_tape
    from ramp_b200.testing import SyntheticPursuitDataset, pursuit_cloud_2d, synthetic_state_dict
    from ramp_b200.unet import TemporalUnetInference
    B, H, k = 16, 64, 2
    net = TemporalUnetInference(n_support_points=H, state_dim=4, max_rows=2 * B)
    m = DynamicGaussianDiffusionModel(model=net, n_diffusion_steps=100, predict_epsilon=True)
    m.load_state_dict(synthetic_state_dict(m, 0))
    m = m.to(cuda).eval()
    ds = SyntheticPursuitDataset()
    cloud = pursuit_cloud_2d(ds).to(cuda)
    static_pts = cloud.reshape(-1, 2)
    np.random.seed(5)
    tape = draw_disc_tape(8, 64)
    s = torch.linspace(0, 1, H)[:, None]
    plan = torch.cat([-0.8 + 1.6 * s, -0.8 + 1.6 * s, torch.full((H, 2), 0.2)], dim=1).to(cuda)
    start = np.float32([-0.6, -0.9])
    dp = DevicePursuer(start, num_points=64, tape=tape, device=cuda)
    g_dev = ReplanGraph(m, B, H, k, cloud, static_pts, torch.zeros(64, 2), static_pts, pursuer=dp)
    g_host = ReplanGraph(m, B, H, k, cloud, static_pts, torch.zeros(64, 2), static_pts)
    assert int(dp.counter) == 0 and np.array_equal(dp.center.cpu().numpy(), start)   # capture warm-up left no trace
    for g in (g_dev, g_host):
        g.x_clean.copy_(plan)
        g.history.copy_(plan[:k + 1])
    c = start
    compared = 0
    for it in range(4):
        c = O.pursuer_step(it, c, plan[k, :2].cpu().numpy(), dp.velocity.numpy(), 0.8, 0.2)
        disc = O.disc_points(c, 0.1, 64, tape[it]).astype(np.float32)
        g_host.disc_points.copy_(torch.from_numpy(disc))
        g_host.disc_center.copy_(torch.from_numpy(c))
        seed = 1000 + it
        a = _replay_seeded(g_host, seed)
        reproducible = torch.equal(a, _replay_seeded(g_host, seed))   # Philox state of a replay follows the reseeded generator
        b = _replay_seeded(g_dev, seed)
        assert (g_dev.disc_center.cpu() - torch.from_numpy(c)).abs().max().item() <= (it + 1) * ULP
        assert (g_dev.disc_points.cpu() - torch.from_numpy(disc)).abs().max().item() <= (it + 2) * ULP
        assert torch.isfinite(b).all()
        if reproducible:      # same noise, clouds equal to an ulp -> same batch to tolerance and the same winner
            compared += 1
            assert (b - a).abs().max().item() < 1e-4
            assert int(g_dev.sel[0]) == int(g_host.sel[0])
    assert int(dp.counter) == 4
    print("ticks compared under identical noise:", compared)


@pytest.mark.gpu
def test_run_inference_with_device_pursuer_in_the_environment(cuda):
    """The planner's own loop (run_inference) with a DevicePursuer standing in for the environment's moving sphere: same
    protocol (`centers`, `radii`, `update_centers`), the pursuer cloud is produced on the device each tick."""
    from ramp_b200.dynamic import DynamicGaussianDiffusionModel
    from ramp_b200.pursuit import DevicePursuer
    from ramp_b200.sample_functions import ddpm_sample_fn
    from ramp_b200.testing import SyntheticPursuitDataset, pursuit_cloud_2d, synthetic_state_dict
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=64, state_dim=4, max_rows=64)
    m = DynamicGaussianDiffusionModel(model=net, n_diffusion_steps=100, predict_epsilon=True)
    m.load_state_dict(synthetic_state_dict(m, 0))
    m = m.to(cuda).eval()
    m.max_iteration = 3
    ds = SyntheticPursuitDataset(device=cuda)
    np.random.seed(0)
    torch.manual_seed(0)
    start = ds.env.obj_extra_list[0].fields[0].centers[0].cpu().numpy()
    dp = DevicePursuer(start, radius=0.1, num_points=64, device=cuda)
    ds.env.obj_extra_list[0].fields[0] = dp
    cloud = pursuit_cloud_2d(ds).to(cuda)
    hc = {0: torch.tensor([-0.8, -0.8, 0.0, 0.0], device=cuda), 63: torch.tensor([0.8, 0.8, 0.0, 0.0], device=cuda)}
    chain, chain_obs, chain_start = m.run_inference({'dataset': ds}, hc, n_samples=6, horizon=64, return_chain=True,
                                                    traj_normalized=torch.zeros(64, 4, device=cuda), obstacle_pts=cloud,
                                                    sample_fn=ddpm_sample_fn)
    assert chain.shape == (4, 1, 64, 4) and len(chain_obs) == 3 and torch.isfinite(chain).all()
    track = torch.cat([c.cpu() for c in chain_obs])
    assert track.shape == (3, 2) and not torch.equal(track[0], track[1]) and not torch.equal(track[1], track[2])
    # each tick moves the pursuer 0.1 * (0.5 / sqrt 2) * |0.8 unit + 0.2 wobble| <= 0.0354
    steps = (track[1:] - track[:-1]).norm(dim=1)
    assert (steps > 0.01).all() and (steps <= 0.0354 * 1.01).all()
    assert int(dp.counter) == 3                                   # one disc cloud per tick, drawn from the tape
