"""TemporalUnetInference — drop-in for mpd/models/diffusion_models/UnetInference.py:40-229.

Same constructor, attributes, parameter names/shapes (reference checkpoints load with
``load_state_dict``) and call signature.  The arithmetic — forward sweep AND the analytic
input-gradient that the reference obtains from autograd (UnetInference.py:19-37) — runs in
hand-written sm_100a kernels behind the C ABI (include/ramp_b200.h: ramp_unet_*).  The nn
sub-modules below are parameter containers only; their ``forward`` is never used on the hot path.
Scene encoders are a cached precompute (UnetInference.py:146-156) and stay in PyTorch.
"""
import ctypes

import torch
import torch.nn as nn

from ._lib import RampError, check, lib, ptr
from .scene_encoders import ObstacleEncoder, ObstacleEncoderSet

UNET_DIM_MULTS = {0: (1, 2, 4), 1: (1, 2, 4, 8)}


class _Holder(nn.Module):
    """Parameter container whose forward must never run (the CUDA engine does the math)."""

    def forward(self, *a, **k):
        raise RampError("parameter container: the score network runs in libramp_b200.so, not in PyTorch")


def _conv_block(cin, cout):
    m = _Holder()
    m.block = nn.Sequential(nn.Conv1d(cin, cout, 5, padding=2), nn.Identity(), nn.GroupNorm(8, cout), nn.Identity(),
                            nn.Identity())
    return m


def _res_block(cin, cout, cond_dim):
    m = _Holder()
    m.blocks = nn.ModuleList([_conv_block(cin, cout), _conv_block(cout, cout)])
    m.cond_mlp = nn.Sequential(nn.Identity(), nn.Linear(cond_dim, cout), nn.Identity())
    m.residual_conv = nn.Conv1d(cin, cout, 1) if cin != cout else nn.Identity()
    return m


def _attention(query_dim, context_dim, inner):
    m = _Holder()
    m.to_q = nn.Linear(query_dim, inner, bias=False)
    m.to_k = nn.Linear(context_dim, inner, bias=False)
    m.to_v = nn.Linear(context_dim, inner, bias=False)
    m.to_out = nn.Sequential(nn.Linear(inner, query_dim), nn.Identity())
    return m


def _transformer_block(dim, context_dim):
    m = _Holder()
    m.attn1 = _attention(dim, dim, dim)
    ff = _Holder()
    proj = _Holder()
    proj.proj = nn.Linear(dim, dim * 8)            # GEGLU: 256 -> 2 x 1024
    ff.net = nn.Sequential(proj, nn.Identity(), nn.Linear(dim * 4, dim))
    m.ff = ff
    m.attn2 = _attention(dim, context_dim, dim)
    m.norm1, m.norm2, m.norm3 = nn.LayerNorm(dim), nn.LayerNorm(dim), nn.LayerNorm(dim)
    return m


def _spatial_transformer(channels, inner, context_dim, depth):
    m = _Holder()
    m.norm = nn.GroupNorm(8, channels, eps=1e-6)
    m.proj_in = nn.Conv1d(channels, inner, 1)
    m.transformer_blocks = nn.ModuleList([_transformer_block(inner, context_dim) for _ in range(depth)])
    m.proj_out = nn.Conv1d(inner, channels, 1)
    nn.init.zeros_(m.proj_out.weight)              # zero_module, layers_attention_mini.py:185
    nn.init.zeros_(m.proj_out.bias)
    return m


def _sampler(kind, dim):
    m = _Holder()
    m.conv = nn.Conv1d(dim, dim, 3, stride=2, padding=1) if kind == "down" else \
        nn.ConvTranspose1d(dim, dim, 4, stride=2, padding=1)
    return m


class TemporalUnetInference(nn.Module):
    def __init__(self, n_support_points=None, state_dim=None, unet_input_dim=32, dim_mults=(1, 2, 4, 8), time_emb_dim=32,
                 self_attention=False, conditioning_embed_dim=4, conditioning_type='attention', attention_num_heads=4,
                 attention_dim_head=64, obstacle_3d=False, precision=3, max_rows=4096, max_scenes=8, **kwargs):
        super().__init__()
        if tuple(dim_mults) != (1, 2, 4, 8) or unet_input_dim != 32 or time_emb_dim != 32 or self_attention \
                or conditioning_type != 'attention' or attention_num_heads != 4 or attention_dim_head != 64:
            raise NotImplementedError("ramp_b200 builds the configuration the inference scripts use: "
                                      "unet_input_dim=32, dim_mults=(1,2,4,8), attention conditioning, 4 heads x 64")
        self.state_dim = state_dim
        self.n_support_points = n_support_points
        self.energy_mode = True
        self.training = False
        self.drop_concept = True
        self.concept_drop_prob = 0.2
        self.cached_scene_latents = None
        self.cached_batch_size = None
        self.enable_caching = True
        self.scene_encoder = ObstacleEncoder() if obstacle_3d else ObstacleEncoderSet()
        self.context_dim = 256 if obstacle_3d else 320
        self.cfg_batch = not obstacle_3d
        self.conditioning_type = conditioning_type
        self.depth_attn = 2
        self.precision, self.max_rows, self.max_scenes = precision, max_rows, max_scenes
        self.max_rows_cap = max_rows     # rows per engine call never exceed this; larger batches are chunked

        dims = [state_dim] + [unet_input_dim * m for m in dim_mults]
        in_out = list(zip(dims[:-1], dims[1:]))
        inner = attention_num_heads * attention_dim_head
        te = _Holder()
        te.encoder = nn.Sequential(nn.Identity(), nn.Linear(32, 128), nn.Identity(), nn.Linear(128, time_emb_dim))
        self.time_mlp = te
        self.downs = nn.ModuleList()
        for i, (ci, co) in enumerate(in_out):
            last = i >= len(in_out) - 1
            self.downs.append(nn.ModuleList([
                _res_block(ci, co, time_emb_dim), _res_block(co, co, time_emb_dim), nn.Identity(),
                _spatial_transformer(co, inner, self.context_dim, self.depth_attn),
                nn.Identity() if last else _sampler("down", co)]))
        mid = dims[-1]
        self.mid_block1 = _res_block(mid, mid, time_emb_dim)
        self.mid_attn = nn.Identity()
        self.mid_attention = _spatial_transformer(mid, inner, self.context_dim, self.depth_attn)
        self.mid_block2 = _res_block(mid, mid, time_emb_dim)
        self.ups = nn.ModuleList()
        for ci, co in reversed(in_out[1:]):
            self.ups.append(nn.ModuleList([
                _res_block(co * 2, ci, time_emb_dim), _res_block(ci, ci, time_emb_dim), nn.Identity(),
                _spatial_transformer(ci, inner, self.context_dim, self.depth_attn), _sampler("up", ci)]))
        fc0 = _conv_block(unet_input_dim, unet_input_dim)
        self.final_conv = nn.Sequential(fc0, nn.Conv1d(unet_input_dim, state_dim, 1))

        self._engine = None          # C handle, built lazily from the current parameters
        self._engine_key = None
        self._scene_key = None
        self._row2scene = None

    # ------------------------------------------------------------------ engine plumbing
    def _unet_state(self):
        return {k: v for k, v in self.state_dict().items() if not k.startswith("scene_encoder.")}

    def _device(self):
        return self.final_conv[1].weight.device

    def release_engine(self):
        if self._engine is not None:
            lib().ramp_unet_destroy(self._engine)
        self._engine, self._engine_key, self._scene_key = None, None, None

    def __del__(self):
        try:
            self.release_engine()
        except Exception:
            pass

    def prepare(self, horizon=None, max_rows=None, max_scenes=None):
        """(Re)build the CUDA engine: upload + repack the weights, size the workspace."""
        dev = self._device()
        if dev.type != "cuda":
            raise RampError("ramp_b200 runs on a CUDA device only (model is on %s); there is no CPU path" % dev)
        horizon = horizon or self.n_support_points
        self.max_rows = max_rows or self.max_rows
        self.max_scenes = max_scenes or self.max_scenes
        self.release_engine()

        class Cfg(ctypes.Structure):
            _fields_ = [(n, ctypes.c_int) for n in ("state_dim", "horizon", "ctx_dim", "precision", "max_rows", "max_scenes")]

        cfg = Cfg(self.state_dim, horizon, self.context_dim, self.precision, self.max_rows, self.max_scenes)
        h = ctypes.c_void_p()
        with torch.cuda.device(dev):
            check(lib().ramp_unet_create(ctypes.byref(cfg), ctypes.byref(h)), "ramp_unet_create")
            self._engine = h
            for name, t in self._unet_state().items():
                t = t.detach().to(device=dev, dtype=torch.float32).contiguous()
                check(lib().ramp_unet_set_param(h, name.encode(), ptr(t), ctypes.c_size_t(t.numel())),
                      "ramp_unet_set_param(%s)" % name)
            check(lib().ramp_unet_finalize(h, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)),
                  "ramp_unet_finalize")
        self._engine_key = (horizon, self.max_rows, self.max_scenes, self.precision, str(dev))
        self._scene_key = None

    def load_state_dict(self, *a, **k):
        out = super().load_state_dict(*a, **k)
        self.release_engine()          # weights changed: repack on next use
        return out

    def _apply(self, fn, *a, **k):
        out = super()._apply(fn, *a, **k)
        if getattr(self, "_engine", None) is not None:
            self.release_engine()      # device / dtype move
        return out

    def workspace_bytes(self):
        return int(lib().ramp_unet_workspace_bytes(self._engine)) if self._engine is not None else 0

    def set_scene_latents(self, latents):
        """latents [S, ctx_dim] (row of zeros = unconditional).  Collapses every cross-attention
        (single context token => softmax == 1) into per-scene constants on the device."""
        lat = latents.detach().to(device=self._device(), dtype=torch.float32).contiguous()
        if self._engine is None:
            self.prepare(max_scenes=max(self.max_scenes, lat.shape[0]))
        if lat.shape[0] > self.max_scenes:
            self.prepare(max_scenes=lat.shape[0])
        lib().ramp_unet_set_scenes.argtypes = None
        check(lib().ramp_unet_set_scenes(self._engine, ptr(lat), int(lat.shape[0]),
                                         ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "ramp_unet_set_scenes")
        self._scene_key = object()

    def energy_grad_rows(self, x, t, row2scene, out=None, want_f=False, x_repeat=1):
        """eps for rows x [N,H,D] at scalar timestep t, row i conditioned on scene row2scene[i]."""
        if self._engine is None or self._scene_key is None:
            raise RampError("call set_scene_latents() first")
        x = x.detach().to(torch.float32).contiguous()
        N = x.shape[0] * x_repeat
        eps = torch.empty((N,) + tuple(x.shape[1:]), device=x.device) if out is None else out
        f = torch.empty_like(eps) if want_f else None
        step = self.max_rows - self.max_rows % x_repeat
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        for s in range(0, N, step):
            n = min(step, N - s)
            check(lib().ramp_unet_energy_grad(self._engine, ptr(x[s // x_repeat:]), n, x_repeat, ctypes.c_float(float(t)),
                                              ptr(row2scene[s:s + n]), ptr(eps[s:s + n]),
                                              ptr(f[s:s + n]) if want_f else ctypes.c_void_p(0), st),
                  "ramp_unet_energy_grad")
        return (eps, f) if want_f else eps

    # ------------------------------------------------------------------ reference call surface
    def cache_scene_encoding(self, obstacle_pts):
        """UnetInference.py:146-156 — recompute only when the batch size changes."""
        was_training = self.scene_encoder.training
        self.scene_encoder.eval()
        try:
            with torch.no_grad():
                if not self.enable_caching:
                    return self.scene_encoder(obstacle_pts)
                b = obstacle_pts.shape[0]
                if self.cached_scene_latents is None or self.cached_batch_size != b:
                    # rows that share a cloud share a latent: encode distinct scenes once
                    uniq, inverse = torch.unique(obstacle_pts.reshape(b, -1), dim=0, return_inverse=True)
                    lat = self.scene_encoder(uniq.reshape(-1, *obstacle_pts.shape[1:]))
                    self.cached_scene_latents = lat[inverse]
                    self.cached_batch_size = b
                    self._scene_key = None
                return self.cached_scene_latents
        finally:
            self.scene_encoder.train(was_training)

    def reset_cache(self):
        self.cached_scene_latents = None
        self.cached_batch_size = None
        self._scene_key = None

    def _masked_latents(self, obstacle_pts, compose):
        """UnetInference.py:186-197 (the in-place masks of the reference also persist in its cache)."""
        lat = self.cache_scene_encoding(obstacle_pts)
        if compose:
            lat[2::3] = 0
        elif self.cfg_batch:
            lat = lat.clone()
            lat[1::2] = 0
        else:
            lat[1, :] = 0.
        return lat

    def _rows(self, x, time, obstacle_pts, compose, want_f):
        if obstacle_pts is None:
            raise RampError("obstacle_pts is required (the reference dereferences it unconditionally)")
        if time.numel() > 1 and not bool((time == time.flatten()[0]).all()):
            raise RampError("all rows of a call must share one timestep (as in every reference sampler)")
        lat = self._masked_latents(obstacle_pts.to(x.dtype), compose)
        uniq, inverse = torch.unique(lat, dim=0, return_inverse=True)
        if self._engine is None or self._engine_key[0] != x.shape[1] or uniq.shape[0] > self.max_scenes:
            self.prepare(horizon=x.shape[1], max_scenes=max(self.max_scenes, uniq.shape[0]))
        self.set_scene_latents(uniq)
        return self.energy_grad_rows(x, float(time.flatten()[0]), inverse.to(torch.int32).contiguous(), want_f=want_f)

    def forward(self, x, time, context=None, x_start=None, obstacle_pts=None, forward_t=None, compose=False):
        if self.energy_mode:
            return self._rows(x, time, obstacle_pts, compose, want_f=False)
        return self.forward_no_energy(x, time, x_start=x_start, obstacle_pts=obstacle_pts, forward_t=forward_t,
                                      compose=compose)

    def forward_no_energy(self, x, time, x_start=None, obstacle_pts=None, forward_t=None, compose=False):
        return self._rows(x, time, obstacle_pts, compose, want_f=True)[1]

    def profile_dump(self, path):
        check(lib().ramp_unet_profile_dump(self._engine, str(path).encode()), "ramp_unet_profile_dump")

    def profile(self, enable, fetch=False):
        """Per-kernel-class CUDA-event timing of the engine (ramp_unet_profile).  With fetch=True returns
        ({class: ms}, {class: launches}) accumulated since the last fetch."""
        if self._engine is None:
            raise RampError("engine not built yet")
        names = ("gemm", "norm", "attention", "elementwise")
        if fetch:
            ms = (ctypes.c_float * 4)()
            cnt = (ctypes.c_int * 4)()
            check(lib().ramp_unet_profile(self._engine, int(bool(enable)), ms, cnt), "ramp_unet_profile")
            return dict(zip(names, list(ms))), dict(zip(names, list(cnt)))
        check(lib().ramp_unet_profile(self._engine, int(bool(enable)), None, None), "ramp_unet_profile")
        return None
