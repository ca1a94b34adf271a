"""north_star's second parity criterion: "identical collision-free / success verdicts on a fixed seed set".

tests/golden/verdicts.pt (oracle/make_golden_verdicts.py) holds, for six noise seeds × six trajectories, what the UNMODIFIED
reference sampled (25 DDPM steps, BASELINE config-1 shape) and what its own metrics / cost code decided about each final
trajectory: box-test collision intensity <= 0.01 (scripts/inference/core/metrics.py:48-126), point-cloud collision at
thresholds 0.02 / 0.05 and the best-of-N index (mpd/models/diffusion_models/cost.py:25-88).

The first reverse step amplifies eps 4602x before the clamp, so a free run from x_T can flip isolated
clamped entries; like tests/test_sampler_gpu.py the strict comparison therefore starts from the reference's state after that
step (x1) and runs the remaining 24 steps freely: EVERY verdict identical and EVERY trajectory's waypoints within
north_star's 1e-3 (FREE_RUN_CAP).  A verdict is exempt
only if the reference's own verdict on its own trajectory flips when the boxes / the distance threshold move by 2e-3
(counted and bounded; the disagreements found are reported all at once).  The full
run from x_T is compared as well and its disagreements are bounded."""
import os

import numpy as np
import pytest
import torch

G = os.path.join(os.path.dirname(__file__), "golden", "verdicts.pt")
MARGIN = 2e-3
# north_star's waypoint tolerance, on the FREE-RUNNING 24 steps, for every trajectory.  Round 1's bf16 + bf16 operand planes
# measured 4.4e-4 median / 1.04e-3 max here; the fp16 + fp16 planes with a loss-scaled backward sweep (csrc/common.cuh)
# sit at the fp32 noise floor (profiles/r2_*).
FREE_RUN_CAP = 1e-3
DEVIATIONS = {}


@pytest.fixture(scope="module")
def gold():
    return torch.load(G, weights_only=False)


def _noise(seed, k, B=6):
    """Draw k of the tape: x_T is draw 0, reverse step j (t = 24 - j) consumes draw 1 + j."""
    return torch.randn((B, 64, 4), generator=torch.Generator("cpu").manual_seed(seed + k), dtype=torch.float32)


def _box_robust(final, centres, sizes):
    """True where the reference's own box verdict does not change when every box grows or shrinks by MARGIN per face."""
    p = final[:, :, None, :2].double()

    def free(pad):
        lo, hi = (centres - sizes[:, None] / 2 - pad).double(), (centres + sizes[:, None] / 2 + pad).double()
        return ~((p >= lo) & (p <= hi)).all(dim=-1).any(dim=-1).any(dim=-1)

    return free(MARGIN) == free(-MARGIN)


def _cloud_robust(final, cloud, thr):
    """True where the reference's own point-cloud verdict is the same at thr - MARGIN and thr + MARGIN."""
    d = torch.cdist(final[:, :, :2].double().reshape(-1, 2), cloud.reshape(-1, 2).double()).reshape(final.shape[0], -1)
    dmin = d.amin(dim=1)
    return (dmin < thr - MARGIN) == (dmin < thr + MARGIN)


def _check_case(case, final, free, hit, best_index, label, problems):
    """final [6, 64, 4] cpu; free / hit[thr] bool [6]; best_index int or None.  Appends every disagreement to `problems`
    (so one run reports all of them) and returns the number of borderline exemptions."""
    dev = (final - case["final"]).abs().amax(dim=(1, 2))
    DEVIATIONS.setdefault(label, []).extend(dev.tolist())
    if dev.max().item() > FREE_RUN_CAP:
        problems.append((label, case["seed"], "waypoints", dev.tolist()))
    exempt = 0
    safe = _box_robust(case["final"], case["centres"], case["sizes"])
    # the intensity verdict counts waypoints (<= 0.01 means none of 64 inside): compare where no waypoint is borderline
    if not torch.equal(free[safe], case["free"][safe]):
        problems.append((label, case["seed"], "box verdict", free.tolist(), case["free"].tolist()))
    exempt += int((~safe).sum())
    for thr in (0.02, 0.05):
        safe = _cloud_robust(case["final"], case["cloud"], thr)
        if not torch.equal(hit[thr][safe], case["cloud_hit"][thr][safe]):
            problems.append((label, case["seed"], "cloud verdict", thr, hit[thr].tolist(), case["cloud_hit"][thr].tolist()))
        exempt += int((~safe).sum())
    if _cloud_robust(case["final"], case["cloud"], 0.02).all() and best_index != case["best_index"]:
        problems.append((label, case["seed"], "best-of-N", best_index, case["best_index"]))
    return exempt


# ------------------------------------------------------------------------------------------------ CPU: the oracle
def test_oracle_verdicts_match_reference_on_two_seeds(gold):
    from oracle import ramp_oracle as O
    from ramp_b200.diffusion import StaticGaussianDiffusionModel
    from ramp_b200.testing import synthetic_state_dict
    from ramp_b200.unet import TemporalUnetInference
    m = StaticGaussianDiffusionModel(model=TemporalUnetInference(n_support_points=64, state_dim=4), n_diffusion_steps=25,
                                     predict_epsilon=True)
    sd, sch = synthetic_state_dict(m, 0), O.make_schedule(25)
    hc = {0: torch.tensor([-0.8, -0.8, 0.0, 0.0]), 63: torch.tensor([0.8, 0.8, 0.0, 0.0])}
    assert sum(int(c["free"].sum()) for c in gold["cases"]) >= 3 and sum(int((~c["free"]).sum()) for c in gold["cases"]) >= 3
    problems = []
    for case in (gold["cases"][0], gold["cases"][3]):
        lat = O.scene_encoder_2d(sd, case["cloud"][None])
        x = case["x1"].clone()
        for j in range(1, 25):
            t = 24 - j
            x = O.ddpm_step(sch, x, t, O.eps_combined(sd, x, t, lat, [2.0]), _noise(case["seed"], 1 + j), 0.5)
            x = O.inpaint(x, hc)
        free = O.metric_collision_intensity(x, case["centres"], case["sizes"]) <= 0.01
        hit = {thr: O.collision_mask(x, case["cloud"].reshape(-1, 2), thr) for thr in (0.02, 0.05)}
        best = O.trajectory_costs(x, case["cloud"].reshape(-1, 2), 0.02)
        _check_case(case, x, free, hit, None if best[0] is None else int(best[4]), "oracle", problems)
    assert not problems, problems
    assert max(DEVIATIONS["oracle"]) <= 1e-3


# ------------------------------------------------------------------------------------------------ GPU: the product
def _product(cuda):
    from ramp_b200.diffusion import StaticGaussianDiffusionModel
    from ramp_b200.testing import synthetic_state_dict
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=64, state_dim=4, max_rows=64)
    m = StaticGaussianDiffusionModel(model=net, n_diffusion_steps=25, predict_epsilon=True)
    m.load_state_dict(synthetic_state_dict(m, 0))
    m = m.to(cuda).eval()
    m.ddim = False
    return m


def _product_verdicts(final, case, cuda):
    from ramp_b200.cost import compute_collision_with_pointcloud, compute_trajectory_costs
    from ramp_b200.metrics import Metrics
    M = Metrics()
    ci = M.compute_collision_intensity(final, case["centres"].to(cuda), case["sizes"].to(cuda))
    met = M.trajectory_success_and_metrics(final, ci, threshold=0.01)
    cloud = case["cloud"].to(cuda)
    hit = {thr: compute_collision_with_pointcloud(final, cloud, collision_threshold=thr).cpu() for thr in (0.02, 0.05)}
    best = compute_trajectory_costs(final, cloud, collision_threshold=0.02)
    return (ci <= 0.01).cpu(), met, hit, None if best[0] is None else int(best[4])


@pytest.mark.gpu
def test_product_verdicts_identical_on_the_fixed_seed_set(cuda, gold):
    from ramp_b200 import ops
    m = _product(cuda)
    hc = {0: torch.tensor([-0.8, -0.8, 0.0, 0.0], device=cuda), 63: torch.tensor([0.8, 0.8, 0.0, 0.0], device=cuda)}
    pins = ops.Pins.from_hard_conds(m._repeat_hard_conds(hc, 6), 6, 64, 4, cuda)
    exempt, n, problems = 0, 0, []
    for case in gold["cases"]:
        cloud = case["cloud"].to(cuda).unsqueeze(0)
        x = case["x1"].to(cuda).contiguous()
        for j in range(1, 25):
            t = 24 - j
            x = m._ddpm_step(x, t, cloud, _noise(case["seed"], 1 + j).to(cuda) if t > 0 else None, 0.5, pins=pins)
        free, met, hit, best = _product_verdicts(x, case, cuda)
        exempt += _check_case(case, x.cpu(), free, hit, best, "cuda", problems)
        if (met["success"], met["n_free_trajectories"]) != (case["success"], case["n_free"]):
            problems.append(("cuda", case["seed"], "success", met["success"], met["n_free_trajectories"], case["success"], case["n_free"]))
        n += 6
    dev = torch.tensor(DEVIATIONS["cuda"][-n:])
    print("verdicts compared on %d trajectories x 3 tests, %d borderline exemptions; free-run waypoint deviation: max %.2e, "
          "median %.2e, %d of %d within 1e-3" % (n, exempt, dev.max(), dev.median(), int((dev <= 1e-3).sum()), n))
    print("per-trajectory deviations:", ["%.1e" % v for v in dev.tolist()])
    assert not problems, problems
    assert exempt == 0                            # no verdict of this fixture is borderline
    assert dev.max().item() <= FREE_RUN_CAP         # all 36, not a quantile


@pytest.mark.gpu
def test_product_verdicts_full_run_from_xT(cuda, gold):
    """run_inference from x_T with the reference's noise tape: trajectories whose clamped first step did not flip (final
    waypoints within 1e-3) must carry identical verdicts; flipped ones are bounded in number."""
    import sys
    sys.path.insert(0, os.path.dirname(__file__))
    from test_sampler_gpu import Tape
    from ramp_b200.sample_functions import ddpm_sample_fn
    m = _product(cuda)
    hc = {0: torch.tensor([-0.8, -0.8, 0.0, 0.0], device=cuda), 63: torch.tensor([0.8, 0.8, 0.0, 0.0], device=cuda)}
    flipped, total, disagree = 0, 0, 0
    for case in gold["cases"]:
        with Tape(case["seed"]):
            chain = m.run_inference(None, hc, n_samples=6, horizon=64, return_chain=True,
                                    traj_normalized=torch.zeros(64, 4, device=cuda), obstacle_pts=case["cloud"].to(cuda),
                                    sample_fn=ddpm_sample_fn, noise_std_extra_schedule_fn=lambda t: 0.5)
        final = chain[-1].contiguous()
        free, _, hit, _ = _product_verdicts(final, case, cuda)
        close = (final.cpu() - case["final"]).abs().amax(dim=(1, 2)) <= 1e-3
        safe = close & _box_robust(case["final"], case["centres"], case["sizes"])
        assert torch.equal(free[safe], case["free"][safe]), case["seed"]
        safe = close & _cloud_robust(case["final"], case["cloud"], 0.02)
        assert torch.equal(hit[0.02][safe], case["cloud_hit"][0.02][safe]), case["seed"]
        flipped += int((~close).sum())
        disagree += int((free != case["free"]).sum())
        total += 6
    print("full runs: %d of %d trajectories deviate > 1e-3 after the clamped first step; %d box verdicts differ"
          % (flipped, total, disagree))
    assert flipped <= total // 4 and disagree <= flipped
