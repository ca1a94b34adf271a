"""Deterministic synthetic weights / inputs (no checkpoints or datasets exist offline).

``synthetic_state_dict`` fills every tensor of a module's state dict from a generator seeded by
the tensor's NAME, so the same values are reproduced on any machine and in any construction
order — the reference model in the build container, the oracle and the CUDA path all load the
same numbers.  ``SpatialTransformer.proj_out`` is zero-initialised in the reference
(layers_attention_mini.py:185); here it gets small random values so parity tests see the
transformer branch."""
import hashlib

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
