"""Scene encoder kernels (csrc/scene_enc.cu, through ramp_scene2d_encode) vs the CPU oracle and the reference golden."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")


def _net(cuda):
    from ramp_b200.testing import synthetic_state_dict
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=64, state_dim=4, max_rows=64)
    sd = synthetic_state_dict(net, seed=0)
    net.load_state_dict(sd)
    return net.to(cuda).eval(), {"model." + k: v for k, v in sd.items()}


@pytest.mark.parametrize("S", [1, 7])
def test_encoder2d_matches_oracle(cuda, S):
    from oracle import ramp_oracle as O
    from ramp_b200.testing import synthetic_cloud_2d
    net, sd = _net(cuda)
    clouds = synthetic_cloud_2d(seed=5, groups=6)                                   # [6, 6, 64, 2]
    g = torch.Generator("cpu").manual_seed(3)
    extra = torch.rand(1, 6, 64, 2, generator=g) * 2 - 1                            # an unclustered cloud
    pts = torch.cat([clouds, extra])[:S].contiguous()
    out = net.scene_encoder(pts.to(cuda)).cpu()
    ref = O.scene_encoder_2d(sd, pts)
    err = (out - ref).abs().max().item()
    print("encoder2d S=%d: max err %.2e (|lat| %.3f)" % (S, err, ref.abs().max()))
    assert out.shape == (S, 320) and err < 2e-5 * max(1.0, ref.abs().max().item())


def test_encoder2d_matches_reference_golden(cuda):
    """The latent the unmodified reference module produced for the golden cloud (tests/golden/unet_rows_static.pt,
    written by oracle/make_golden.py from /root/reference)."""
    from test_sampler_gpu import build
    g = torch.load(os.path.join(G, "unet_rows_static.pt"), weights_only=False)
    m, _ = build(cuda, "static", 25)
    out = m.model.scene_encoder(g["cloud"][None].to(cuda)).cpu()[0]
    err = (out - g["latent"].reshape(-1)).abs().max().item()
    print("encoder2d vs reference golden: max err %.2e" % err)
    assert err < 2e-5 * max(1.0, g["latent"].abs().max().item())


def test_encoder2d_is_batch_invariant_and_reproducible(cuda):
    from ramp_b200.testing import synthetic_cloud_2d
    net, _ = _net(cuda)
    pts = synthetic_cloud_2d(seed=9, groups=6).to(cuda)
    full = net.scene_encoder(pts)
    assert torch.equal(full, net.scene_encoder(pts))
    for i in (0, 5):
        assert torch.equal(net.scene_encoder(pts[i:i + 1].contiguous())[0], full[i])
    assert net.scene_encoder(pts[:0].contiguous()).shape == (0, 320)


def test_encoder2d_rejects_cpu_and_other_shapes(cuda):
    from ramp_b200._lib import RampError
    net, _ = _net(cuda)
    with pytest.raises(RampError):
        net.scene_encoder(torch.zeros(1, 6, 64, 2))
    with pytest.raises(RampError):
        net.scene_encoder(torch.zeros(1, 5, 64, 2, device=cuda))


def _net3d(cuda):
    from ramp_b200.testing import synthetic_state_dict
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=64, state_dim=6, obstacle_3d=True, max_rows=64)
    sd = synthetic_state_dict(net, seed=0)
    net.load_state_dict(sd)
    return net.to(cuda).eval(), {"model." + k: v for k, v in sd.items()}


@pytest.mark.parametrize("S,n_obs,n_pts", [(1, 25, 64), (3, 10, 37), (2, 1, 1), (2, 32, 130)])
def test_encoder3d_matches_oracle(cuda, S, n_obs, n_pts):
    from oracle import ramp_oracle as O
    net, sd = _net3d(cuda)
    g = torch.Generator("cpu").manual_seed(4)
    pts = torch.rand(S, n_obs, n_pts, 3, generator=g) * 2 - 1
    out = net.scene_encoder(pts.to(cuda)).cpu()
    ref = O.scene_encoder_3d(sd, pts)
    err = (out - ref).abs().max().item()
    print("encoder3d S=%d n_obs=%d n_pts=%d: max err %.2e (|lat| %.3f)" % (S, n_obs, n_pts, err, ref.abs().max()))
    assert out.shape == (S, 256) and err < 2e-5 * max(1.0, ref.abs().max().item())


def test_encoder3d_matches_reference_golden(cuda):
    """The latent of the unmodified reference ObstacleEncoder for the golden cloud (tests/golden/unet_rows_3d.pt)."""
    from test_sampler_gpu import build
    g = torch.load(os.path.join(G, "unet_rows_3d.pt"), weights_only=False)
    m, _ = build(cuda, "3d", 25)
    out = m.model.scene_encoder(g["cloud"][None].to(cuda)).cpu()[0]
    err = (out - g["latent"].reshape(-1)).abs().max().item()
    print("encoder3d vs reference golden: max err %.2e" % err)
    assert err < 2e-5 * max(1.0, g["latent"].abs().max().item())


def test_encoder3d_limits(cuda):
    from ramp_b200._lib import RampError
    net, _ = _net3d(cuda)
    with pytest.raises(RampError):
        net.scene_encoder(torch.zeros(1, 33, 8, 3, device=cuda))
    with pytest.raises(RampError):
        net.scene_encoder(torch.zeros(1, 4, 8, 3))
    assert net.scene_encoder(torch.zeros(0, 4, 8, 3, device=cuda)).shape == (0, 256)
