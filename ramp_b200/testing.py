"""Deterministic synthetic weights / inputs (no checkpoints or datasets exist offline).

``synthetic_state_dict`` fills every tensor of a module's state dict from a generator seeded by
the tensor's NAME, so the same values are reproduced on any machine and in any construction
order — the reference model in the build container, the oracle and the CUDA path all load the
same numbers.  ``SpatialTransformer.proj_out`` is zero-initialised in the reference
(layers_attention_mini.py:185); here it gets small random values so parity tests see the
transformer branch."""
import hashlib

import numpy as np

import torch


def _seed(name, seed):
    return int.from_bytes(hashlib.sha256(("%d/%s" % (seed, name)).encode()).digest()[:6], "little")


def synthetic_state_dict(module_or_shapes, seed=0, skip=("betas", "alphas", "sqrt_", "log_one", "posterior_")):
    sd = module_or_shapes.state_dict() if hasattr(module_or_shapes, "state_dict") else module_or_shapes
    out = {}
    for name, ref in sd.items():
        shape = tuple(ref.shape) if hasattr(ref, "shape") else tuple(ref)
        if any(name.startswith(s) for s in skip) or name.endswith("num_batches_tracked") or name.endswith("div_term"):
            out[name] = ref.clone() if hasattr(ref, "clone") else None
            continue
        g = torch.Generator("cpu").manual_seed(_seed(name, seed))
        leaf = name.rsplit(".", 1)[-1]
        if name.endswith("running_var"):
            t = 0.8 + 0.4 * torch.rand(shape, generator=g)
        elif name.endswith("running_mean"):
            t = 0.1 * torch.randn(shape, generator=g)
        elif len(shape) == 1 and leaf == "weight":            # norm scales
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif len(shape) == 1:                                  # biases (incl. in_proj_bias)
            t = 0.05 * torch.randn(shape, generator=g)
        else:
            fan_in = 1
            for s in shape[1:]:
                fan_in *= s
            bound = (1.0 / fan_in) ** 0.5
            if ".proj_out." in name:
                bound *= 0.5
            t = (torch.rand(shape, generator=g) * 2 - 1) * bound
        out[name] = t.float()
    return out


def synthetic_cloud_2d(seed=0, n_obs=6, n_pts=64, groups=None):
    """box centres U(-0.7,0.7)^2, 64 points U(centre +- 0.13) each."""
    g = torch.Generator("cpu").manual_seed(1000 + seed)
    shape = (n_obs,) if groups is None else (groups, n_obs)
    centres = (torch.rand(*shape, 1, 2, generator=g) * 2 - 1) * 0.7
    return (centres + (torch.rand(*shape, n_pts, 2, generator=g) * 2 - 1) * 0.13).float()


def synthetic_cloud_3d(seed=0, n_obs=25, n_pts=64):
    g = torch.Generator("cpu").manual_seed(2000 + seed)
    centres = (torch.rand(n_obs, 1, 3, generator=g) * 2 - 1) * 0.8
    return (centres + (torch.rand(n_obs, n_pts, 3, generator=g) * 2 - 1) * 0.1).float()


def tape(seed, shapes):
    """Noise tape: draw k = randn(shape_k) from Generator(seed + k) on CPU (same rule as oracle/ref_harness.NoiseTape)."""
    out = []
    for k, shape in enumerate(shapes):
        g = torch.Generator("cpu").manual_seed(seed + k)
        out.append(torch.randn(tuple(shape), generator=g, dtype=torch.float32))
    return out


# ---- synthetic pursuit environment (the attributes DynamicGaussianDiffusionModel reads from context['dataset'].env,
#      diffusion_model_dynamic.py:515-518,396-397) -------------------------------------------------------------------
class _BoxField:
    def __init__(self, centers, sizes):
        self.centers, self.sizes = centers, sizes


class _PursuerField:
    """Stand-in for torch_robotics MultiSphereFieldDynamics: `centers` [1, 2], `radii` [1], `update_centers(t, pos)`
    moves the pursuer `speed` towards the first row of `pos` plus a small t-dependent swirl (deterministic)."""

    def __init__(self, center, radius=0.1, speed=0.04):
        self.centers = torch.tensor([list(center)], dtype=torch.float32)
        self.radii = torch.tensor([radius], dtype=torch.float32)
        self.speed = speed

    def update_centers(self, t, current_state=None):
        if t is None or current_state is None:
            return
        c = self.centers[0].double().cpu()
        p = torch.as_tensor(current_state).detach().double().cpu().reshape(-1, 2)[0]
        d = p - c
        n = float(d.norm())
        if n > 0:
            d = d / n
        swirl = torch.tensor([np.sin(2 * np.pi * 0.1 * t), np.cos(2 * np.pi * 0.1 * t)], dtype=torch.float64)
        new = (c + self.speed * (0.8 * d + 0.2 * swirl)).clamp(-1, 1)
        self.centers = new.float().reshape(1, 2).to(self.centers.device)


class _Obj:
    def __init__(self, field):
        self.fields = [field]


class _Env:
    def __init__(self, boxes, pursuer):
        self.obj_fixed_list = [_Obj(boxes)]
        self.obj_extra_list = [_Obj(pursuer)]


class SyntheticPursuitDataset:
    """context['dataset'] for the dynamic model: 4 static boxes just outside the [-1, 1]^2 workspace walls (inner faces at 1.18) (random-init
    weights give noise-like plans that fill the workspace: obstacles inside it would leave no collision-free
    trajectory, which the reference cannot handle, diffusion_model_dynamic.py:548) + one pursuer 0.47 away from the start."""

    def __init__(self, pursuer_start=(-0.35, -0.95), pursuer_speed=0.03, device="cpu"):
        centers = torch.tensor([[1.28, -0.6], [-1.28, 0.5], [0.1, 1.28], [-0.15, -1.28]], dtype=torch.float32, device=device)
        sizes = torch.tensor([[0.2, 0.5], [0.2, 0.6], [0.6, 0.2], [0.5, 0.2]], dtype=torch.float32, device=device)
        self.env = _Env(_BoxField(centers, sizes), _PursuerField(pursuer_start, speed=pursuer_speed))


def pursuit_cloud_2d(dataset, n_pts=64, seed=0):
    """[6, 64, 2] conditioning / cost cloud: points on the four boxes plus two repeats (inference_dynamic.py:142)."""
    g = torch.Generator("cpu").manual_seed(3000 + seed)
    f = dataset.env.obj_fixed_list[0].fields[0]
    c, s = f.centers.cpu(), f.sizes.cpu()
    pts = c[:, None, :] + (torch.rand(4, n_pts, 2, generator=g) - 0.5) * s[:, None, :]
    return torch.cat([pts, pts[[1, 3]]], dim=0).float()
