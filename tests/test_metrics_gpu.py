"""Device metrics (csrc/metrics_ops.cu through ramp_b200.metrics.Metrics) vs the reference golden and the CPU oracle."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")


def _same(a, b, tol):
    if b is None or (isinstance(b, float) and np.isnan(b)):
        return a is None or np.isnan(a)
    return abs(a - b) < tol * max(1.0, abs(b))


def test_metrics_match_reference_golden(cuda):
    from ramp_b200.metrics import Metrics
    g = torch.load(os.path.join(G, "metrics.pt"), weights_only=False)
    M = Metrics()
    for c in g["cases"]:
        x = c["x"].to(cuda)
        ci = M.compute_collision_intensity(x, c["centers"].to(cuda), c["sizes"].to(cuda))
        assert torch.equal(ci.cpu(), c["ci"])
        assert (M.compute_path_length(x).cpu() - c["path_length"]).abs().max() < 1e-5
        assert (M.compute_smoothness(x).cpu() - c["smoothness"]).abs().max() < 1e-4
        assert (M.compute_smoothness(x, x[..., 2:].contiguous()).cpu() - c["smoothness"]).abs().max() < 1e-4
        m, r = M.trajectory_success_and_metrics(x, ci, threshold=0.01), c["metrics"]
        assert m["success"] == r["success"] and m["n_free_trajectories"] == r["n_free_trajectories"]
        assert m["free_trajectories"].shape[0] == r["n_free_trajectories"]
        assert _same(m["collision_intensity"], r["collision_intensity"], 1e-6)
        for k in ("path_length", "path_length_std", "waypoint_variance"):
            assert _same(m[k], r[k], 2e-5), (k, m[k], r[k])
        if c["variance_all"] is not None:
            assert _same(float(M.compute_variance_waypoints(x)), float(c["variance_all"]), 2e-5)


def test_waypoint_variance_full_batch_vs_oracle(cuda):
    """n = 1024 trajectories (the headline batch): 64 x 1024^2 pair distances vs the fp64 restatement, bit-reproducible."""
    from oracle import ramp_oracle as O
    from ramp_b200.metrics import Metrics
    g = torch.Generator("cpu").manual_seed(6)
    x = torch.cat([torch.cumsum(torch.randn(1024, 64, 2, generator=g) * 0.03, dim=1), torch.randn(1024, 64, 2, generator=g)], -1)
    a = Metrics.compute_variance_waypoints(x.to(cuda))
    assert torch.equal(a, Metrics.compute_variance_waypoints(x.to(cuda)))
    ref = sum(float(O.metric_variance_waypoints(x[:, h:h + 1])) for h in range(64))      # one step at a time: 8 MB each
    print("waypoint variance n=1024: %.7f vs oracle %.7f" % (float(a), ref))
    assert abs(float(a) - ref) < 1e-5 * ref


def test_metric_edge_cases(cuda):
    from ramp_b200._lib import RampError
    from ramp_b200.metrics import Metrics
    M = Metrics()
    x = torch.zeros(3, 64, 4, device=cuda)
    assert float(M.compute_variance_waypoints(x)) == 0.0                 # identical trajectories
    assert float(M.compute_variance_waypoints(x[:1])) == 0.0             # fewer than two
    assert M.compute_path_length(x[:0]).item() == 0.0                    # metrics.py:36-37
    ci = M.compute_collision_intensity(x, torch.zeros(1, 2, device=cuda), torch.tensor([0.0], device=cuda))
    assert torch.equal(ci.cpu(), torch.ones(3))                          # inclusive bounds: the point sits on a zero-size box
    none = M.trajectory_success_and_metrics(x, ci, threshold=0.01)
    assert none["success"] == 0 and none["n_free_trajectories"] == 0 and none["path_length"] is None
    with pytest.raises(RampError):
        M.compute_path_length(torch.zeros(3, 64, 4))
