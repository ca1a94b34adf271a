"""Reference harness — TEST INFRASTRUCTURE ONLY (never imported by the product path).

Imports the unmodified reference (``/root/reference``) on CPU, with ``matplotlib`` stubbed,
and provides (a) a *noise tape* that replaces ``torch.randn`` / ``torch.randn_like`` so the
reference and the CUDA path consume identical x_T and per-step noise, and (b) seeded model
factories with re-randomised ``SpatialTransformer.proj_out`` (zero-initialised in the
reference, mpd/models/layers/layers_attention_mini.py:185 — otherwise 92 % of the network is
invisible to parity).

This file only works where ``/root/reference`` exists (the build container). It is used to
validate ``oracle/ramp_oracle.py`` and to generate ``tests/golden/*.pt``
(``oracle/make_golden.py``).  Nothing on the GPU box imports it.
"""
import contextlib
import os
import sys
import types

import torch

_TORCH_RANDN = torch.randn  # bound before any patching


def _default_root():
    """/root/reference in the build container; oracle/_ref (oracle/build_ref.py's untracked copy) on the GPU box."""
    if os.path.isdir("/root/reference/mpd"):
        return "/root/reference"
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


REF_ROOT = os.environ.get("RAMP_REFERENCE_ROOT") or _default_root()


def reference_available():
    return os.path.isdir(os.path.join(REF_ROOT, "mpd"))


def _stub_matplotlib():
    if "matplotlib" in sys.modules:
        return

    class _Stub(types.ModuleType):
        def __getattr__(self, name):
            if name.startswith("__"):
                raise AttributeError(name)
            return _Stub(self.__name__ + "." + name)

        def __call__(self, *a, **k):
            return _Stub(self.__name__ + "()")

    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.animation", "matplotlib.patches",
                 "matplotlib.collections", "matplotlib.colors", "matplotlib.cm"):
        sys.modules[name] = _Stub(name)


_REF = None


def import_reference():
    """Return the reference's ``mpd.models`` namespace."""
    global _REF
    if _REF is None:
        if not reference_available():
            raise RuntimeError("reference tree not present at %s" % REF_ROOT)
        _stub_matplotlib()
        for pth in (os.path.join(REF_ROOT, "deps", "torch_robotics"), REF_ROOT):
            if pth not in sys.path:
                sys.path.insert(0, pth)
        with contextlib.redirect_stdout(open(os.devnull, "w")):
            import mpd.models as m  # noqa
        _REF = m
    return _REF


class NoiseTape:
    """Deterministic replacement for torch.randn / torch.randn_like.

    Draw k of the tape is ``torch.randn(shape, generator=Generator(seed + k))`` on CPU in fp32,
    moved to the requested device.  Call order in the reference: x_T at
    diffusion_model_static.py:240 / :359, then one ``randn_like`` per ``ddpm_sample_fn``
    (sample_functions.py:39), ``q_sample`` noise (diffusion_model_dynamic.py:672).
    """

    def __init__(self, seed):
        self.seed = int(seed)
        self.k = 0
        self.draws = []

    def draw(self, shape, device=None, dtype=torch.float32):
        g = torch.Generator("cpu").manual_seed(self.seed + self.k)
        self.k += 1
        out = _TORCH_RANDN(tuple(shape), generator=g, dtype=torch.float32)
        self.draws.append(out)
        return out.to(device=device, dtype=dtype)

    @contextlib.contextmanager
    def patched(self):
        orig_randn, orig_like = torch.randn, torch.randn_like

        def randn(*size, device=None, dtype=None, **kw):
            if len(size) == 1 and isinstance(size[0], (tuple, list, torch.Size)):
                size = tuple(size[0])
            return self.draw(size, device=device, dtype=dtype or torch.get_default_dtype())

        def randn_like(x, **kw):
            return self.draw(x.shape, device=x.device, dtype=x.dtype)

        torch.randn, torch.randn_like = randn, randn_like
        try:
            yield self
        finally:
            torch.randn, torch.randn_like = orig_randn, orig_like


def randomise_zero_modules(unet, seed=1, std=0.02):
    """Give every zero-initialised ``proj_out`` (and BatchNorm running stats) non-trivial values."""
    g = torch.Generator("cpu").manual_seed(seed)
    with torch.no_grad():
        for name, p in unet.named_parameters():
            if ".proj_out." in name:
                p.copy_(torch.randn(p.shape, generator=g) * std)
        for name, b in unet.named_buffers():
            if name.endswith("running_mean"):
                b.copy_(torch.randn(b.shape, generator=g) * 0.1)
            elif name.endswith("running_var"):
                b.copy_(1.0 + 0.2 * torch.rand(b.shape, generator=g))


def make_reference_model(kind="static", n_diffusion_steps=25, compose=False, use_apf=False, seed=0,
                         horizon=64, dtype=torch.float32):
    """Random-init reference model (CPU).  kind in {static, dynamic, 3d}."""
    m = import_reference()
    torch.manual_seed(seed)
    state_dim = 6 if kind == "3d" else 4
    with contextlib.redirect_stdout(open(os.devnull, "w")):
        unet = m.TemporalUnetInference(state_dim=state_dim, n_support_points=horizon, unet_input_dim=32,
                                       dim_mults=m.UNET_DIM_MULTS[1], obstacle_3d=(kind == "3d"))
    randomise_zero_modules(unet)
    cls = {"static": m.StaticGaussianDiffusionModel, "dynamic": m.DynamicGaussianDiffusionModel,
           "3d": m.GaussianDiffusionModel3d}[kind]
    model = cls(model=unet, variance_schedule="exponential", n_diffusion_steps=n_diffusion_steps,
                predict_epsilon=True, compose=compose, use_apf=use_apf, training=False)
    model = model.to(dtype).eval()
    for p in model.parameters():
        p.requires_grad_(False)
    return model


def synthetic_cloud_2d(seed=0, n_obs=6, n_pts=64, groups=None):
    """BASELINE config-1 cloud: box centres U(-0.7,0.7)^2, points U(centre +- 0.13)."""
    g = torch.Generator("cpu").manual_seed(1000 + seed)
    shape = (n_obs,) if groups is None else (groups, n_obs)
    centres = (torch.rand(*shape, 1, 2, generator=g) * 2 - 1) * 0.7
    pts = centres + (torch.rand(*shape, n_pts, 2, generator=g) * 2 - 1) * 0.13
    return pts.float()


def synthetic_cloud_3d(seed=0, n_obs=25, n_pts=64):
    g = torch.Generator("cpu").manual_seed(2000 + seed)
    centres = (torch.rand(n_obs, 1, 3, generator=g) * 2 - 1) * 0.8
    pts = centres + (torch.rand(n_obs, n_pts, 3, generator=g) * 2 - 1) * 0.1
    return pts.float()


def hard_conds_2d(horizon=64):
    return {0: torch.tensor([-0.8, -0.8, 0.0, 0.0]), horizon - 1: torch.tensor([0.8, 0.8, 0.0, 0.0])}


def hard_conds_3d(horizon=64):
    return {0: torch.tensor([-0.8, -0.8, -0.8, 0.0, 0.0, 0.0]),
            horizon - 1: torch.tensor([0.8, 0.8, 0.8, 0.0, 0.0, 0.0])}
