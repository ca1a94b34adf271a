"""Score-network energy gradient: CUDA engine (C ABI) vs the CPU oracle on identical weights/inputs."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(cuda, dim3, precision=3, H=64):
    from ramp_b200.unet import TemporalUnetInference
    from ramp_b200.testing import synthetic_state_dict
    D = 6 if dim3 else 4
    net = TemporalUnetInference(n_support_points=H, state_dim=D, obstacle_3d=dim3, precision=precision, max_rows=64)
    sd = synthetic_state_dict(net, seed=0)
    net.load_state_dict(sd)
    return net.to(cuda), {"model." + k: v for k, v in sd.items()}, D


@pytest.mark.parametrize("dim3,H,N", [(False, 64, 6), (True, 64, 5), (False, 48, 3), (False, 64, 37)])
def test_energy_grad_matches_oracle(cuda, dim3, H, N):
    from oracle import ramp_oracle as O
    net, sd, D = _setup(cuda, dim3, H=H)
    g = torch.Generator("cpu").manual_seed(11)
    x = torch.randn(N, H, D, generator=g)
    lat = torch.randn(3, net.context_dim, generator=g) * 0.5
    lat[2] = 0
    r2s = torch.arange(N) % 3
    t = 13
    net.prepare(horizon=H, max_rows=64, max_scenes=4)
    net.set_scene_latents(lat.to(cuda))
    eps, f = net.energy_grad_rows(x.to(cuda), t, r2s.to(torch.int32).to(cuda), want_f=True)
    torch.cuda.synchronize()
    f_ref = O.unet_forward(sd, x, torch.full((N,), t), lat[r2s])
    e_ref = O.energy_grad(sd, x, torch.full((N,), t), lat[r2s])
    ferr = (f.cpu() - f_ref).abs().max().item()
    eerr = (eps.cpu() - e_ref).abs().max().item()
    print("f err %.3e (|f| %.3f)  eps err %.3e (|eps| %.3f)" % (ferr, f_ref.abs().max(), eerr, e_ref.abs().max()))
    assert ferr < 2e-4 * max(1.0, f_ref.abs().max().item())
    assert eerr < 2e-4 * max(1.0, e_ref.abs().max().item())


@pytest.mark.parametrize("dim3", [False, True])
def test_energy_grad_is_bit_reproducible(cuda, dim3):
    """Every reduction of the engine runs in a fixed order (no float atomics): the same inputs give the same bits."""
    net, _, D = _setup(cuda, dim3)
    g = torch.Generator("cpu").manual_seed(5)
    x = torch.randn(33, 64, D, generator=g).to(cuda)
    lat = (torch.randn(3, net.context_dim, generator=g) * 0.5).to(cuda)
    r2s = (torch.arange(33) % 3).to(torch.int32).to(cuda)
    net.prepare(horizon=64, max_rows=64, max_scenes=4)
    net.set_scene_latents(lat)
    first = net.energy_grad_rows(x, 7, r2s).clone()
    for _ in range(3):
        again = net.energy_grad_rows(x, 7, r2s)
        assert torch.equal(first, again)
