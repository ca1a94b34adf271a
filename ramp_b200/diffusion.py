"""GaussianDiffusionModel drop-ins: StaticGaussianDiffusionModel, GaussianDiffusionModel3d,
DynamicGaussianDiffusionModel (mpd/models/diffusion_models/diffusion_model_{static,3d,dynamic}.py).

Same constructors, registered buffers (state-dict keys), attributes and `run_inference` /
`conditional_sample` / `p_sample_loop` / `ddim_p_sample_loop` / `warmup` call surface.  The host code
only sequences kernels: per step ONE score-network call (ramp_unet_energy_grad over B*(G+1) rows,
trajectories broadcast on the device) and ONE fused step kernel (ramp_step); the APF and the
cost/selection are device kernels too (no KD-tree, no host round trip).
"""
from copy import copy

import numpy as np
import torch
import torch.nn as nn

from . import ops
from ._lib import RampError
from .sample_functions import apply_hard_conditioning, ddpm_sample_fn, extract  # noqa: F401 (re-exported)


_TORCH_RANDN, _TORCH_RANDN_LIKE = torch.randn, torch.randn_like     # to notice a patched RNG (a noise tape)


def make_timesteps(batch_size, i, device):
    return torch.full((batch_size,), i, device=device, dtype=torch.long)


# ---- variance schedules (mpd/models/diffusion_models/helpers.py:20-46) --------------------------
def cosine_beta_schedule(n_diffusion_steps, s=0.008, a_min=0, a_max=0.999, dtype=torch.float32):
    steps = n_diffusion_steps + 1
    x = np.linspace(0, steps, steps)
    ac = np.cos(((x / steps) + s) / (1 + s) * np.pi * 0.5) ** 2
    ac = ac / ac[0]
    betas = 1 - (ac[1:] / ac[:-1])
    return torch.tensor(np.clip(betas, a_min=a_min, a_max=a_max), dtype=dtype)


def exponential_beta_schedule(n_diffusion_steps, beta_start=1e-4, beta_end=1.0):
    x = torch.linspace(0, n_diffusion_steps, n_diffusion_steps)
    b0, b1 = torch.tensor(beta_start), torch.tensor(beta_end)
    a = 1 / n_diffusion_steps * torch.log(b1 / b0)
    return b0 * torch.exp(a * x)


class _LossStub(nn.Module):
    """Keeps the `loss_fn` attribute (training is out of scope); contributes no state-dict keys."""

    def __init__(self):
        super().__init__()
        self.register_buffer('weights', None)


class _DiffusionBase(nn.Module):
    # guidance constants hard-coded in the reference samplers
    cfg_weight = 2.0
    compose_weights = (2.0, 2.0)

    def __init__(self, model=None, variance_schedule='exponential', n_diffusion_steps=100, clip_denoised=True,
                 predict_epsilon=False, loss_type='l2', context_model=None, compose=False, use_apf=False,
                 training=False, **kwargs):
        super().__init__()
        self.model = model
        self.context_model = context_model
        self.n_diffusion_steps = n_diffusion_steps
        self.compose = compose
        self.APF = use_apf
        self.energy_mode = True
        self.training = training
        self.state_dim = self.model.state_dim
        if not clip_denoised:
            raise RampError("clip_denoised=False is rejected by the reference too (diffusion_model_static.py:170)")
        if variance_schedule == 'cosine':
            betas = cosine_beta_schedule(n_diffusion_steps, s=0.008, a_min=0, a_max=0.999)
        elif variance_schedule == 'exponential':
            betas = exponential_beta_schedule(n_diffusion_steps, beta_start=1e-4, beta_end=1.0)
        else:
            raise NotImplementedError
        alphas = 1. - betas
        ac = torch.cumprod(alphas, axis=0)
        ac_prev = torch.cat([torch.ones(1), ac[:-1]])
        self.clip_denoised = clip_denoised
        self.predict_epsilon = predict_epsilon
        reg = self.register_buffer
        reg('betas', betas)
        reg('alphas_cumprod', ac)
        reg('alphas_cumprod_prev', ac_prev)
        reg('sqrt_alphas_cumprod', torch.sqrt(ac))
        reg('sqrt_one_minus_alphas_cumprod', torch.sqrt(1. - ac))
        reg('log_one_minus_alphas_cumprod', torch.log(1. - ac))
        reg('sqrt_recip_alphas_cumprod', torch.sqrt(1. / ac))
        reg('sqrt_recipm1_alphas_cumprod', torch.sqrt(1. / ac - 1))
        post_var = betas * (1. - ac_prev) / (1. - ac)
        reg('posterior_variance', post_var)
        reg('posterior_log_variance_clipped', torch.log(torch.clamp(post_var, min=1e-20)))
        reg('posterior_mean_coef1', betas * np.sqrt(ac_prev) / (1. - ac))
        reg('posterior_mean_coef2', (1. - ac_prev) * np.sqrt(alphas) / (1. - ac))
        self.final_alpha_cumprod = torch.tensor([1.0, ])
        self.loss_fn = _LossStub()
        self._sched_host = None
        self._scene_sig = None
        self._r2s = {}

    # ------------------------------------------------------------------ schedule scalars (host copies)
    def _apply(self, fn, *a, **k):
        out = super()._apply(fn, *a, **k)
        self._sched_host = None
        self._r2s = {}
        return out

    def load_state_dict(self, *a, **k):
        out = super().load_state_dict(*a, **k)
        self._sched_host = None
        self._scene_sig = None
        return out

    def _sched(self):
        """fp32 per-timestep scalars, computed with the same torch fp32 ops as the reference and kept on
        the host so that no step needs a device->host sync (the reference syncs at sample_functions.py:25)."""
        if self._sched_host is None:
            g = lambda n: getattr(self, n).detach().float().cpu()
            ac = g('alphas_cumprod')
            self._sched_host = {
                'sqrt_recip': g('sqrt_recip_alphas_cumprod'), 'sqrt_recipm1': g('sqrt_recipm1_alphas_cumprod'),
                'coef1': g('posterior_mean_coef1'), 'coef2': g('posterior_mean_coef2'),
                'sigma': torch.exp(0.5 * g('posterior_log_variance_clipped')),
                'ac': ac, 'sqrt_ac': g('sqrt_alphas_cumprod'), 'sqrt_1mac': g('sqrt_one_minus_alphas_cumprod'),
                'ac_pow': ac ** 0.5, 'beta_pow': (1 - ac) ** 0.5,
            }
        return self._sched_host

    def _ddim_scalars(self, t, t_prev):
        """diffusion_model_static.py:266-276,321-332 with eta = 0."""
        s = self._sched()
        a_prev = s['ac'][t_prev] if t_prev >= 0 else torch.tensor(1.0)
        return dict(sa_t=float(s['ac_pow'][t]), sb_t=float(s['beta_pow'][t]), sa_prev=float(a_prev ** 0.5),
                    sdir=float((1 - a_prev - torch.tensor(0.0) ** 2) ** 0.5))

    # ------------------------------------------------------------------ conditioning
    def _set_scenes(self, groups):
        """groups [G, n_obs, n_pts, d]: one conditional cloud per group; an all-zero latent is appended
        as the unconditional branch (UnetInference.py:190-197).  Encoded once per distinct input."""
        sig = (groups.data_ptr(), tuple(groups.shape), groups._version, self.model._engine_key)
        if sig == self._scene_sig and self.model._scene_key is not None:
            return self._G
        unet = self.model
        was = unet.scene_encoder.training
        unet.scene_encoder.eval()
        with torch.no_grad():
            lat = unet.scene_encoder(groups.to(torch.float32))
        unet.scene_encoder.train(was)
        lat = torch.cat([lat, torch.zeros(1, lat.shape[1], device=lat.device)], dim=0)
        unet.set_scene_latents(lat)
        self._G = groups.shape[0]
        self._scene_sig = (groups.data_ptr(), tuple(groups.shape), groups._version, self.model._engine_key)
        self._scene_src = groups          # pointer-keyed caches hold their key tensor: its memory cannot be recycled under them
        return self._G

    def _ensure_engine(self, horizon, rows):
        unet = self.model
        key = unet._engine_key
        need = min(rows, unet.max_rows_cap)
        if key is None or key[0] != horizon or key[1] < need:
            unet.prepare(horizon=horizon, max_rows=need if key is None else max(need, key[1]))
            self._scene_sig = None

    def _eps_rows(self, x, t, G):
        """One network call over B*(G+1) rows, row b*(G+1)+g = trajectory b under scene g."""
        B = x.shape[0]
        key = (B, G, x.device)
        if key not in self._r2s:
            self._r2s = {key: self._row_scenes(B, G, x.device)}
        self._eps_strides = (None, 1)                      # row of (b, g) = b*(G+1) + g
        return self.model.energy_grad_rows(x, t, self._r2s[key], x_repeat=G + 1)

    def _row_scenes(self, B, G, device):
        """Scene (latent row) of network row b*(G+1)+g: group g for g < G, the zero latent (index G) for g == G."""
        return (torch.arange(B * (G + 1), device=device, dtype=torch.int32) % (G + 1)).contiguous()

    def _groups_and_weights(self, obstacle_pts, compose):
        raise NotImplementedError

    # ------------------------------------------------------------------ one reverse step
    def _ddpm_step(self, x, t, obstacle_pts, noise, noise_scale, compose=False, pins=None, out=None, forward_t=None):
        groups, w = self._groups_and_weights(obstacle_pts, compose)
        self._ensure_engine(x.shape[1], x.shape[0] * (groups.shape[0] + 1))
        G = self._set_scenes(groups)
        x = x.contiguous()
        eps = self._eps_rows(x, t, G)
        s = self._sched()
        out = torch.empty_like(x) if out is None else out
        return ops.step(x, eps, G, w, out, 0, pins=pins, eps_bstride=self._eps_strides[0], eps_gstride=self._eps_strides[1],
                        sqrt_recip=s['sqrt_recip'][t], sqrt_recipm1=s['sqrt_recipm1'][t],
                        coef1=s['coef1'][t], coef2=s['coef2'][t], sigma=s['sigma'][t], noise_scale=noise_scale,
                        noise=noise, predict_epsilon=self.predict_epsilon)

    def _x0_step(self, x, t, obstacle_pts, compose=False, x0_out=None, eps_out=None):
        groups, w = self._groups_and_weights(obstacle_pts, compose)
        self._ensure_engine(x.shape[1], x.shape[0] * (groups.shape[0] + 1))
        G = self._set_scenes(groups)
        x = x.contiguous()
        eps = self._eps_rows(x, t, G)
        s = self._sched()
        x0_out = torch.empty_like(x) if x0_out is None else x0_out
        ops.step(x, eps, G, w, None, 2, eps_bstride=self._eps_strides[0], eps_gstride=self._eps_strides[1],
                        sqrt_recip=s['sqrt_recip'][t], sqrt_recipm1=s['sqrt_recipm1'][t], x0_out=x0_out,
                 eps_out=eps_out, predict_epsilon=self.predict_epsilon)
        return x0_out

    def _ddim_step(self, x, t, t_prev, obstacle_pts, compose=False, pins=None, out=None):
        groups, w = self._groups_and_weights(obstacle_pts, compose)
        self._ensure_engine(x.shape[1], x.shape[0] * (groups.shape[0] + 1))
        G = self._set_scenes(groups)
        x = x.contiguous()
        eps = self._eps_rows(x, t, G)
        s = self._sched()
        out = torch.empty_like(x) if out is None else out
        return ops.step(x, eps, G, w, out, 1, pins=pins, eps_bstride=self._eps_strides[0], eps_gstride=self._eps_strides[1],
                        sqrt_recip=s['sqrt_recip'][t], sqrt_recipm1=s['sqrt_recipm1'][t],
                        predict_epsilon=self.predict_epsilon, **self._ddim_scalars(t, t_prev))

    # ------------------------------------------------------------------ CUDA-graph fast path of the DDPM loop
    # 'auto': the third call with the same shapes captures the whole loop (graphs.DdpmLoopGraph) and later calls replay it —
    # one graph launch instead of ~7,500 host launches per sampling (north_star (3)); True: capture at the first call;
    # False: always eager.  A replay draws from torch's default CUDA generator exactly like the eager loop (bit-identical
    # under the same torch.manual_seed, tests/test_loops_gpu.py); a patched torch.randn (noise tape) disables the path.
    cuda_graph = 'auto'

    def _loop_graph(self, shape, hard_conds, obstacle_pts, sample_fn, noise_fn, extra_steps):
        if self.cuda_graph is False or sample_fn is not ddpm_sample_fn or not self.betas.is_cuda or self.APF:
            return None        # (the APF branch packs its cloud outside the graph: eager only)
        if torch.randn is not _TORCH_RANDN or torch.randn_like is not _TORCH_RANDN_LIKE:
            return None
        if torch.cuda.is_current_stream_capturing():
            return None
        B, H, D = shape
        steps = list(reversed(range(-extra_steps, self.n_diffusion_steps)))
        scales = tuple(1.0 if noise_fn is None else float(noise_fn(i)) for i in steps)
        graphs = self.__dict__.setdefault('_loop_graphs', {})
        key = (B, H, D, scales, bool(self.compose), bool(self.APF), tuple(obstacle_pts.shape), str(self.betas.device))
        ent = graphs.setdefault(key, {'calls': 0, 'graph': None, 'src': None})
        pins = ops.Pins.from_hard_conds(hard_conds, B, H, D, self.betas.device)
        g = ent['graph']
        if g is not None and (g.pins.val.shape != pins.val.shape or not torch.equal(g.pins.mask, pins.mask)):
            g = ent['graph'] = None
        if g is None:
            ent['calls'] += 1
            if self.cuda_graph is not True and ent['calls'] <= 2:
                return None
            from .graphs import DdpmLoopGraph
            if len(graphs) > 4:                        # each graph pins a chain and noise buffers: keep a handful
                for k in [k for k in graphs if k != key]:
                    del graphs[k]
            # the capture's warm-up passes draw from the default generator: put its state back afterwards, so that THIS call
            # (the first replay) consumes exactly the draws the eager loop would have consumed
            dev = self.betas.device
            rng = torch.cuda.get_rng_state(dev)
            g = ent['graph'] = DdpmLoopGraph(self, B, H, hard_conds, obstacle_pts, noise_scale=list(scales),
                                             extra_steps=extra_steps, own_cloud=True)
            torch.cuda.set_rng_state(rng, dev)
            ent['src'] = None
        g.pins.val.copy_(pins.val)
        src = (obstacle_pts.data_ptr(), obstacle_pts._version)
        if ent['src'] is None or ent['src'][0] != src:
            g.set_cloud(obstacle_pts)
            ent['src'] = (src, obstacle_pts)          # holding the tensor keeps its pointer from being recycled under the key
        return g

    def _replay_loop(self, g):
        """-> chain [T+1, B, H, D] as a fresh tensor (the graph's own chain buffer is overwritten by the next replay), or
        None when the engine was rebuilt since the capture (the caller falls back to the eager loop; a new graph is
        captured on a later call)."""
        try:
            return g.replay().clone()
        except RampError:
            for ent in self.__dict__.get('_loop_graphs', {}).values():
                if ent['graph'] is g:
                    ent['graph'], ent['calls'] = None, 0
            return None

    # ------------------------------------------------------------------ reference helper surface
    def predict_start_from_noise(self, x_t, t, noise):
        if self.predict_epsilon:
            return extract(self.sqrt_recip_alphas_cumprod, t, x_t.shape) * x_t - \
                extract(self.sqrt_recipm1_alphas_cumprod, t, x_t.shape) * noise
        return noise

    def q_posterior(self, x_start, x_t, t):
        mean = extract(self.posterior_mean_coef1, t, x_t.shape) * x_start + \
            extract(self.posterior_mean_coef2, t, x_t.shape) * x_t
        return mean, extract(self.posterior_variance, t, x_t.shape), \
            extract(self.posterior_log_variance_clipped, t, x_t.shape)

    def q_sample(self, x_start, t, noise=None):
        """diffusion_model_dynamic.py:671-680 (single shared timestep)."""
        if noise is None:
            noise = torch.randn_like(x_start)
        s = self._sched()
        ti = int(t.flatten()[0]) if torch.is_tensor(t) else int(t)
        return ops.q_sample(x_start.contiguous(), noise.contiguous(), torch.empty_like(x_start), s['sqrt_ac'][ti],
                            s['sqrt_1mac'][ti])

    def ddim_set_timesteps(self, num_inference_steps):
        self.num_inference_steps = num_inference_steps
        ratio = self.n_diffusion_steps // self.num_inference_steps
        return (np.arange(0, num_inference_steps) * ratio).round()[::-1].copy().astype(np.int64)

    def forward(self, cond, *args, **kwargs):
        raise NotImplementedError

    def _repeat_hard_conds(self, hard_conds, n_samples):
        out = copy(hard_conds)
        for k, v in out.items():
            out[k] = v.unsqueeze(0).expand(n_samples, -1).contiguous() if v.dim() == 1 else v
        return out


class _CloudCache:
    """Packed device copies of APF / cost clouds, keyed on the source tensor."""

    def __init__(self):
        self._d = {}

    def get(self, pts):
        key = (pts.data_ptr(), tuple(pts.shape), pts._version)
        if key not in self._d:
            if len(self._d) > 8:
                self._d.clear()
            self._d[key] = (ops.CloudIndex(pts), pts)      # the key tensor stays alive with its entry
        return self._d[key][0]


class StaticGaussianDiffusionModel(_DiffusionBase):
    """mpd/models/diffusion_models/diffusion_model_static.py:21-512 (sampling side)."""
    cfg_weight = 2.0                 # :164
    compose_weights = (2.0, 2.0)     # :208
    apf_threshold, apf_strength, apf_window, apf_iters, apf_from = 0.07, 0.1, 7, 3, 2   # :298-315

    def __init__(self, model=None, variance_schedule='exponential', n_diffusion_steps=100, clip_denoised=True,
                 predict_epsilon=False, loss_type='l2', context_model=None, compose=False, use_apf=False,
                 training=False, **kwargs):
        super().__init__(model, variance_schedule, n_diffusion_steps, clip_denoised, predict_epsilon, loss_type,
                         context_model, compose, use_apf, training, **kwargs)
        self.ddim_num_inference_steps = 8 if (compose and use_apf) else 5   # :40
        self.ddim = True                                                      # :41
        self._groups_cache = None
        self._clouds = _CloudCache()

    def _groups_and_weights(self, obstacle_pts, compose):
        """CFG: obstacle_pts [1, 6, 64, 2] (or [6, 64, 2]) -> one group, w = 2 (:149-165).
        compose: [G, 6, 64, 2]; the reference asserts G == 2 (:197) with w = (2, 2) (:208); more groups follow the
        commented 3-set template at diffusion_model_3d.py:165-174 with the same per-group weight."""
        key = (obstacle_pts.data_ptr(), tuple(obstacle_pts.shape), obstacle_pts._version, bool(compose))
        if self._groups_cache is None or self._groups_cache[0] != key:
            g = obstacle_pts if obstacle_pts.dim() == 4 else obstacle_pts.unsqueeze(0)
            if not compose:
                g = g[:1]
            self._groups_cache = (key, g.to(torch.float32).contiguous(), obstacle_pts)
        g = self._groups_cache[1]
        G = g.shape[0]
        if not compose:
            return g, [self.cfg_weight]
        w = list(self.compose_weights) + [self.compose_weights[-1]] * max(0, G - len(self.compose_weights))
        return g, w[:G]

    # DDPM step with the potential field on the posterior mean --------------------------------------
    ddpm_apf_threshold, ddpm_apf_strength, ddpm_apf_window, ddpm_apf_after = 0.07, 0.1, 5, 20    # :172-176

    def _ddpm_apf_active(self, compose, forward_t):
        """:172 — only the non-compose p_mean_variance has the branch (p_mean_variance_compose, :188-229, does not)."""
        return self.APF and not compose and forward_t is not None and forward_t > self.ddpm_apf_after

    def _ddpm_step(self, x, t, obstacle_pts, noise, noise_scale, compose=False, pins=None, out=None, forward_t=None):
        """One DDPM step.  With `use_apf` and DDPM sampling the reference pushes the posterior MEAN through `avoidance`
        (threshold 0.07, strength 0.1, window 5) for forward_t > 20, then adds the noise (:172-181, sample_functions.py:32-48);
        here: fused step without noise -> nearest point -> windowed push -> noise + inpainting (ramp_step mode 3)."""
        if not self._ddpm_apf_active(compose, forward_t):
            return super()._ddpm_step(x, t, obstacle_pts, noise, noise_scale, compose=compose, pins=pins, out=out)
        mean = super()._ddpm_step(x, t, obstacle_pts, None, noise_scale, compose=compose, pins=None)
        cloud = self._clouds.get(obstacle_pts)
        dist, idx = cloud.query(mean, self.ddpm_apf_threshold)
        pushed = ops.apf_static(mean, cloud, dist, idx, self.ddpm_apf_threshold, self.ddpm_apf_strength,
                                self.ddpm_apf_window)
        out = torch.empty_like(x) if out is None else out
        s = self._sched()
        return ops.step(pushed, None, 1, [0.0], out, 3, pins=pins, sigma=s['sigma'][t], noise_scale=noise_scale, noise=noise)

    # reference-shaped entry points ------------------------------------------------------------
    def p_mean_variance(self, x, hard_conds, context, t, traj_normalized=None, obstacle_pts=None, forward_t=None,
                        compose=False):
        """:149-186 — returns (model_mean, posterior_variance, posterior_log_variance[, x_recon, e_comb])."""
        ti = int(t[0])
        eps = torch.empty_like(x)
        x0 = self._x0_step(x, ti, obstacle_pts, compose=compose, eps_out=eps)
        mean, var, logvar = self.q_posterior(x_start=x0, x_t=x, t=t)
        if self.ddim:
            return mean, var, logvar, x0, eps
        if self._ddpm_apf_active(compose, forward_t):                      # :172-181
            cloud = self._clouds.get(obstacle_pts)
            dist, idx = cloud.query(mean.contiguous(), self.ddpm_apf_threshold)
            mean = ops.apf_static(mean.contiguous(), cloud, dist, idx, self.ddpm_apf_threshold, self.ddpm_apf_strength,
                                  self.ddpm_apf_window)
        return mean, var, logvar

    def p_mean_variance_compose(self, x, hard_conds, context, t, traj_normalized=None, obstacle_pts=None,
                                forward_t=None, compose=True):
        return self.p_mean_variance(x, hard_conds, context, t, traj_normalized, obstacle_pts, forward_t, compose=True)

    @torch.no_grad()
    def p_sample_loop(self, shape, hard_conds, context=None, return_chain=False, traj_normalized=None,
                      obstacle_pts=None, sample_fn=ddpm_sample_fn, n_diffusion_steps_without_noise=0, **sample_kwargs):
        """:231-256.  The chain is written in place: step k reads chain[k] and writes chain[k+1]."""
        device = self.betas.device
        B, H, D = shape
        if not self.compose and obstacle_pts.dim() == 3:
            obstacle_pts = obstacle_pts.unsqueeze(0)
        noise_fn = sample_kwargs.get('noise_std_extra_schedule_fn')
        g = self._loop_graph(shape, hard_conds, obstacle_pts, sample_fn, noise_fn, n_diffusion_steps_without_noise)
        chain = self._replay_loop(g) if g is not None else None
        if chain is not None:
            return (chain[-1], chain.transpose(0, 1)) if return_chain else chain[-1]
        pins = ops.Pins.from_hard_conds(hard_conds, B, H, D, device)
        n_total = self.n_diffusion_steps + n_diffusion_steps_without_noise
        chain = torch.empty(n_total + 1, B, H, D, device=device)
        chain[0] = torch.randn(shape, device=device)
        ops.inpaint(chain[0], pins)
        k = 0
        for i in reversed(range(-n_diffusion_steps_without_noise, self.n_diffusion_steps)):
            x = chain[k]
            if sample_fn is ddpm_sample_fn:
                ti = max(i, 0)
                noise = torch.randn_like(x)
                ns = 1.0 if noise_fn is None else float(noise_fn(i))
                self._ddpm_step(x, ti, obstacle_pts, noise if ti > 0 else None, ns, compose=self.compose, pins=pins,
                                out=chain[k + 1], forward_t=k)
            else:   # foreign sample_fn: keep the reference protocol
                t = make_timesteps(B, i, device)
                xn, _ = sample_fn(self, x, hard_conds, context, t, traj_normalized=traj_normalized,
                                  obstacle_pts=obstacle_pts, forward_t=k, compose=self.compose, **sample_kwargs)
                chain[k + 1] = ops.inpaint(xn.contiguous(), pins)
            k += 1
        x = chain[-1]
        if return_chain:
            return x, chain.transpose(0, 1)
        return x

    def set_apf_cloud(self, points):
        """Extension (no reference counterpart): use `points` ([P, d] tensor or an `ops.CloudIndex`) as the potential
        field's cloud instead of the conditioning cloud.  The reference derives the APF cloud from `obstacle_pts`
        (:304-311), which the scene encoder hard-wires to 6 x 64 points (obstacle_encoder.py:95); BASELINE's dense clouds
        (4 k ... 1 M points) can only enter this way.  `None` restores the reference behaviour."""
        self._apf_override = None if points is None else (points if isinstance(points, ops.CloudIndex) else ops.CloudIndex(points))

    def _apf_cloud(self, obstacle_pts):
        """:304-311 — compose uses the first six obstacles of set 0 and the first four of set 1."""
        if getattr(self, '_apf_override', None) is not None:
            return self._apf_override
        if self.compose:
            key = (obstacle_pts.data_ptr(), tuple(obstacle_pts.shape), obstacle_pts._version)
            if getattr(self, '_apf_key', None) != key:
                self._apf_pts = torch.cat([obstacle_pts[0], obstacle_pts[1][:4]], dim=0).reshape(-1, 2).contiguous()
                self._apf_key, self._apf_src = key, obstacle_pts
            return self._clouds.get(self._apf_pts)
        return self._clouds.get(obstacle_pts)

    def _apf(self, x0, cloud, pins):
        """:298-319 — three query+push+inpaint rounds on x0 (each re-queries the pushed waypoints)."""
        for _ in range(self.apf_iters):
            dist, idx = cloud.query(x0, self.apf_threshold)
            ops.apf_static(x0, cloud, dist, idx, self.apf_threshold, self.apf_strength, self.apf_window, out=x0)
            ops.inpaint(x0, pins)
        return x0

    def ddim_p_sample(self, x, hard_conds, context, t, obstacle_pts, traj_normalized=None, forward_t=None, eta=0.0,
                      use_clipped_model_output=False, pins=None, out=None):
        """:259-333 (eta must be 0, as in every reference call)."""
        assert use_clipped_model_output and eta == 0.0
        ti = int(t[0]) if torch.is_tensor(t) else int(t)
        t_prev = ti - self.n_diffusion_steps // self.ddim_num_inference_steps
        if pins is None:
            pins = ops.Pins.from_hard_conds(hard_conds, x.shape[0], x.shape[1], x.shape[2], x.device)
        if self.APF and forward_t is not None and forward_t >= self.apf_from:
            x0 = self._x0_step(x, ti, obstacle_pts, compose=self.compose)
            self._apf(x0, self._apf_cloud(obstacle_pts), pins)
            out = torch.empty_like(x) if out is None else out
            return ops.ddim_update(x.contiguous(), x0, out, pins=None, **self._ddim_scalars(ti, t_prev))
        return self._ddim_step(x, ti, t_prev, obstacle_pts, compose=self.compose, pins=None, out=out)

    def ddim_p_sample_loop(self, shape, hard_conds, context=None, return_chain=False, traj_normalized=None,
                           obstacle_pts=None, t_start_guide=float('inf'), guide=None, n_guide_steps=1, **sample_kwargs):
        """:347-384."""
        device = self.betas.device
        B, H, D = shape
        pins = ops.Pins.from_hard_conds(hard_conds, B, H, D, device)
        time_idx = self.ddim_set_timesteps(self.ddim_num_inference_steps)
        if not self.compose and obstacle_pts.dim() == 3:
            obstacle_pts = obstacle_pts.unsqueeze(0)
        chain = torch.empty(len(time_idx) + 1, B, H, D, device=device)
        chain[0] = torch.randn(shape, device=device)
        ops.inpaint(chain[0], pins)
        for fwd, i in enumerate(time_idx):
            self.ddim_p_sample(chain[fwd], hard_conds, context, int(i), obstacle_pts, traj_normalized=traj_normalized,
                               forward_t=fwd, eta=0.0, use_clipped_model_output=True, pins=pins, out=chain[fwd + 1])
            ops.inpaint(chain[fwd + 1], pins)
        if return_chain:
            return chain[-1], chain.transpose(0, 1)
        return chain[-1]

    @torch.no_grad()
    def conditional_sample(self, hard_conds, horizon=None, batch_size=1, ddim=False, traj_normalized=None,
                           obstacle_pts=None, **sample_kwargs):
        """:386-399 — `self.ddim` (True by default, :41) selects the loop, not the `ddim` argument."""
        horizon = horizon or self.model.n_support_points
        shape = (batch_size, horizon, self.state_dim)
        if self.ddim:
            sample_kwargs.pop('sample_fn', None)
            sample_kwargs.pop('noise_std_extra_schedule_fn', None)
            sample_kwargs.pop('n_diffusion_steps_without_noise', None)
            return self.ddim_p_sample_loop(shape, hard_conds, traj_normalized=traj_normalized, obstacle_pts=obstacle_pts,
                                           **sample_kwargs)
        for k in ('guide', 'n_guide_steps', 't_start_guide'):
            sample_kwargs.pop(k, None)
        return self.p_sample_loop(shape, hard_conds, traj_normalized=traj_normalized, obstacle_pts=obstacle_pts,
                                  **sample_kwargs)

    @torch.no_grad()
    def warmup(self, horizon=64, traj_normalized=None, obstacle_pts=None, batch_size=None, device='cuda'):
        """:405-433 — one network evaluation; also builds the engine and fills the scene cache."""
        x = torch.randn((batch_size, horizon, self.state_dim), device=device)
        if not self.compose and obstacle_pts.dim() == 3:
            obstacle_pts = obstacle_pts.unsqueeze(0)
        self._x0_step(x, 1, obstacle_pts.to(device), compose=self.compose)

    @torch.no_grad()
    def run_inference(self, context=None, hard_conds=None, n_samples=1, return_chain=False, traj_normalized=None,
                      obstacle_pts=None, **diffusion_kwargs):
        """:437-463 — returns the chain as [steps+1, n_samples, H, D] (or its last entry)."""
        hard_conds = self._repeat_hard_conds(hard_conds, n_samples)
        context = copy(context)
        _, chain = self.conditional_sample(hard_conds, context=context, batch_size=n_samples, ddim=False,
                                           return_chain=True, traj_normalized=traj_normalized,
                                           obstacle_pts=obstacle_pts, **diffusion_kwargs)
        chain = chain.transpose(0, 1)   # 'b diffsteps h d -> diffsteps b h d'
        return chain if return_chain else chain[-1]


class GaussianDiffusionModel3d(_DiffusionBase):
    """mpd/models/diffusion_models/diffusion_model_3d.py:19-391 (sampling side).

    The reference evaluates exactly one trajectory per call (`out[0]`, `out[1]` at :150); here the batch is
    vectorised: every trajectory is evaluated under [conditional..., unconditional] on B*(G+1) network rows,
    which for B == 1 is the reference computation."""
    cfg_weight = 5.75                # :150
    compose_weights = (5.0, 5.0)     # :169-170

    def __init__(self, model=None, variance_schedule='exponential', n_diffusion_steps=100, clip_denoised=True,
                 predict_epsilon=False, loss_type='l2', context_model=None, compose=False, use_apf=False,
                 training=False, **kwargs):
        super().__init__(model, variance_schedule, n_diffusion_steps, clip_denoised, predict_epsilon, loss_type,
                         context_model, compose, use_apf, training, **kwargs)
        self._groups_cache = None

    def _groups_and_weights(self, obstacle_pts, compose):
        key = (obstacle_pts.data_ptr(), tuple(obstacle_pts.shape), obstacle_pts._version, bool(compose))
        if self._groups_cache is None or self._groups_cache[0] != key:
            g = obstacle_pts if obstacle_pts.dim() == 4 else obstacle_pts.unsqueeze(0)
            if not compose:
                g = g[:1]
            self._groups_cache = (key, g.to(torch.float32).contiguous(), obstacle_pts)
        g = self._groups_cache[1]
        G = g.shape[0]
        if not compose:
            return g, [self.cfg_weight]
        w = list(self.compose_weights) + [self.compose_weights[-1]] * max(0, G - len(self.compose_weights))
        return g, w[:G]

    def p_mean_variance(self, x, hard_conds, context, t, traj_normalized=None, obstacle_pts=None, forward_t=None,
                        compose=False):
        """:147-161."""
        x0 = self._x0_step(x, int(t[0]), obstacle_pts, compose=compose)
        return self.q_posterior(x_start=x0, x_t=x, t=t)

    def p_mean_variance_compose(self, x, hard_conds, context, t, traj_normalized=None, obstacle_pts=None,
                                forward_t=None, compose=True):
        """:163-182."""
        return self.p_mean_variance(x, hard_conds, context, t, traj_normalized, obstacle_pts, forward_t, compose=True)

    @torch.no_grad()
    def p_sample_loop(self, shape, hard_conds, context=None, return_chain=False, traj_normalized=None,
                      obstacle_pts=None, sample_fn=ddpm_sample_fn, n_diffusion_steps_without_noise=0, **sample_kwargs):
        """:184-218 (resample_steps is fixed to 1 there)."""
        device = self.betas.device
        B, H, D = shape
        if not self.compose and obstacle_pts.dim() == 3:
            obstacle_pts = obstacle_pts.unsqueeze(0)
        noise_fn = sample_kwargs.get('noise_std_extra_schedule_fn')
        g = self._loop_graph(shape, hard_conds, obstacle_pts, sample_fn, noise_fn, n_diffusion_steps_without_noise)
        chain = self._replay_loop(g) if g is not None else None
        if chain is not None:
            return (chain[-1], chain.transpose(0, 1)) if return_chain else chain[-1]
        pins = ops.Pins.from_hard_conds(hard_conds, B, H, D, device)
        n_total = self.n_diffusion_steps + n_diffusion_steps_without_noise
        chain = torch.empty(n_total + 1, B, H, D, device=device)
        chain[0] = torch.randn(shape, device=device)
        ops.inpaint(chain[0], pins)
        k = 0
        for i in reversed(range(-n_diffusion_steps_without_noise, self.n_diffusion_steps)):
            x = chain[k]
            if sample_fn is ddpm_sample_fn:
                ti = max(i, 0)
                noise = torch.randn_like(x)
                ns = 1.0 if noise_fn is None else float(noise_fn(i))
                self._ddpm_step(x, ti, obstacle_pts, noise if ti > 0 else None, ns, compose=self.compose, pins=pins,
                                out=chain[k + 1])
            else:
                t = make_timesteps(B, i, device)
                xn, _ = sample_fn(self, x, hard_conds, context, t, traj_normalized=traj_normalized,
                                  obstacle_pts=obstacle_pts, forward_t=k, compose=self.compose, **sample_kwargs)
                chain[k + 1] = ops.inpaint(xn.contiguous(), pins)
            k += 1
        if return_chain:
            return chain[-1], chain.transpose(0, 1)
        return chain[-1]

    @torch.no_grad()
    def conditional_sample(self, hard_conds, horizon=None, batch_size=1, ddim=False, traj_normalized=None,
                           obstacle_pts=None, **sample_kwargs):
        """:289-300."""
        if ddim:
            raise NotImplementedError("GaussianDiffusionModel3d.ddim_sample calls the network without a cloud in the "
                                      "reference (diffusion_model_3d.py:247) and cannot run there either")
        horizon = horizon or self.model.n_support_points
        for k in ('guide', 'n_guide_steps', 't_start_guide'):
            sample_kwargs.pop(k, None)
        return self.p_sample_loop((batch_size, horizon, self.state_dim), hard_conds, traj_normalized=traj_normalized,
                                  obstacle_pts=obstacle_pts, **sample_kwargs)

    @torch.no_grad()
    def warmup(self, horizon=64, device='cuda'):
        """:306-311 — builds the engine (the reference's call passes no cloud and cannot execute)."""
        self.model.prepare(horizon=horizon)

    @torch.no_grad()
    def run_inference(self, context=None, hard_conds=None, n_samples=1, return_chain=False, traj_normalized=None,
                      obstacle_pts=None, **diffusion_kwargs):
        """:313-344."""
        hard_conds = self._repeat_hard_conds(hard_conds, n_samples)
        context = copy(context)
        _, chain = self.conditional_sample(hard_conds, context=context, batch_size=n_samples, return_chain=True,
                                           traj_normalized=traj_normalized, obstacle_pts=obstacle_pts,
                                           **diffusion_kwargs)
        chain = chain.transpose(0, 1)
        return chain if return_chain else chain[-1]
