"""CPU oracle for the RAMP reverse-diffusion hot path — TEST INFRASTRUCTURE ONLY.

A plain-PyTorch (CPU, fp32, autograd) restatement of the reference algorithm for the path
named by BASELINE.json: score network energy gradient, CFG / compositional combine, DDPM /
DDIM step, inpainting, nearest-point APF, trajectory cost + selection.  It consumes a
reference-format ``state_dict`` and externally supplied noise, so that the CUDA path, this
oracle and the reference itself can be compared on identical inputs.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl
reference`` legs may import this module.  The product (``ramp_b200``) never does.

Parity pinning: the reference ships no tests or golden vectors, so this oracle is
pinned against outputs of the reference itself, generated in the build container by
``oracle/make_golden.py`` (imports /root/reference) and committed under ``tests/golden/``;
``tests/test_oracle_golden.py`` re-checks them on every run.

Every function cites the reference file:line it restates (paths relative to the reference
root).
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------------------
# schedules  (mpd/models/diffusion_models/helpers.py:40-46, diffusion_model_static.py:49-88)
# --------------------------------------------------------------------------------------


def exponential_betas(T, beta_start=1e-4, beta_end=1.0):
    x = torch.linspace(0, T, T)
    a = (1.0 / T) * torch.log(torch.tensor(beta_end) / torch.tensor(beta_start))
    return torch.tensor(beta_start) * torch.exp(a * x)


def cosine_betas(T, s=0.008, a_min=0.0, a_max=0.999):
    steps = T + 1
    x = np.linspace(0, steps, steps)
    ac = np.cos(((x / steps) + s) / (1 + s) * np.pi * 0.5) ** 2
    ac = ac / ac[0]
    betas = 1 - (ac[1:] / ac[:-1])
    return torch.tensor(np.clip(betas, a_min, a_max), dtype=torch.float32)


def make_schedule(T, kind="exponential"):
    """All registered buffers of the reference samplers, by their state-dict names."""
    betas = exponential_betas(T) if kind == "exponential" else cosine_betas(T)
    alphas = 1.0 - betas
    ac = torch.cumprod(alphas, 0)
    ac_prev = torch.cat([torch.ones(1), ac[:-1]])
    post_var = betas * (1.0 - ac_prev) / (1.0 - ac)
    return {
        "betas": betas,
        "alphas_cumprod": ac,
        "alphas_cumprod_prev": ac_prev,
        "sqrt_alphas_cumprod": torch.sqrt(ac),
        "sqrt_one_minus_alphas_cumprod": torch.sqrt(1.0 - ac),
        "log_one_minus_alphas_cumprod": torch.log(1.0 - ac),
        "sqrt_recip_alphas_cumprod": torch.sqrt(1.0 / ac),
        "sqrt_recipm1_alphas_cumprod": torch.sqrt(1.0 / ac - 1),
        "posterior_variance": post_var,
        "posterior_log_variance_clipped": torch.log(torch.clamp(post_var, min=1e-20)),
        "posterior_mean_coef1": betas * np.sqrt(ac_prev) / (1.0 - ac),
        "posterior_mean_coef2": (1.0 - ac_prev) * np.sqrt(alphas) / (1.0 - ac),
    }


def ddim_timesteps(T, n):
    """diffusion_model_static.py:336-345."""
    ratio = T // n
    return (np.arange(0, n) * ratio).round()[::-1].copy().astype(np.int64)


# --------------------------------------------------------------------------------------
# score network  (UnetInference.py:157-224, layers.py, layers_attention_mini.py)
# --------------------------------------------------------------------------------------


def _mish(x):
    return x * torch.tanh(F.softplus(x))


def time_embedding(sd, t, prefix="model.time_mlp.encoder."):
    """layers.py:233-259 — sinusoidal(32) -> Linear -> Mish -> Linear.  t: [N] (any int/float)."""
    half = 16
    freq = torch.exp(torch.arange(half, dtype=torch.float32) * -(math.log(10000.0) / (half - 1)))
    e = t.to(torch.float32)[:, None] * freq[None, :]
    e = torch.cat([e.sin(), e.cos()], dim=-1)
    h = F.linear(e, sd[prefix + "1.weight"], sd[prefix + "1.bias"])
    h = _mish(h)
    return F.linear(h, sd[prefix + "3.weight"], sd[prefix + "3.bias"])


def _conv_block(sd, p, x, n_groups=8):
    """layers.py:280-297 — Conv1d(k5,p2) -> GroupNorm(8, eps 1e-5) -> Mish.  x: [N,C,L]."""
    w = sd[p + "block.0.weight"]
    x = F.conv1d(x, w, sd[p + "block.0.bias"], padding=w.shape[-1] // 2)
    x = F.group_norm(x, n_groups, sd[p + "block.2.weight"], sd[p + "block.2.bias"], eps=1e-5)
    return _mish(x)


def _res_block(sd, p, x, t_emb):
    """layers.py:327-361."""
    cond = F.linear(F.silu(t_emb), sd[p + "cond_mlp.1.weight"], sd[p + "cond_mlp.1.bias"])
    h = _conv_block(sd, p + "blocks.0.", x) + cond[:, :, None]
    h = _conv_block(sd, p + "blocks.1.", h)
    if (p + "residual_conv.weight") in sd:
        res = F.conv1d(x, sd[p + "residual_conv.weight"], sd[p + "residual_conv.bias"])
    else:
        res = x
    return h + res


def _attention(sd, p, x, ctx, heads=4):
    """layers_attention_mini.py:83-127.  x: [N,L,256]; ctx: None (self) or [N,1,ctx_dim]."""
    src = x if ctx is None else ctx
    q = F.linear(x, sd[p + "to_q.weight"])
    k = F.linear(src, sd[p + "to_k.weight"])
    v = F.linear(src, sd[p + "to_v.weight"])
    N, L, inner = q.shape
    d = inner // heads

    def split(t):
        return t.reshape(N, -1, heads, d).permute(0, 2, 1, 3)

    q, k, v = split(q), split(k), split(v)
    sim = torch.matmul(q, k.transpose(-1, -2)) * (d ** -0.5)
    attn = sim.softmax(dim=-1)
    out = torch.matmul(attn, v).permute(0, 2, 1, 3).reshape(N, L, inner)
    return F.linear(out, sd[p + "to_out.0.weight"], sd[p + "to_out.0.bias"])


def _transformer_block(sd, p, x, ctx):
    """layers_attention_mini.py:130-149."""
    def ln(name, t):
        return F.layer_norm(t, (t.shape[-1],), sd[p + name + ".weight"], sd[p + name + ".bias"], eps=1e-5)

    x = _attention(sd, p + "attn1.", ln("norm1", x), None) + x
    x = _attention(sd, p + "attn2.", ln("norm2", x), ctx) + x
    u = F.linear(ln("norm3", x), sd[p + "ff.net.0.proj.weight"], sd[p + "ff.net.0.proj.bias"])
    a, gate = u.chunk(2, dim=-1)
    x = F.linear(a * F.gelu(gate), sd[p + "ff.net.2.weight"], sd[p + "ff.net.2.bias"]) + x
    return x


def _spatial_transformer(sd, p, x, ctx, depth=2):
    """layers_attention_mini.py:152-202.  x: [N,C,L]; GroupNorm eps 1e-6."""
    x_in = x
    x = F.group_norm(x, 8, sd[p + "norm.weight"], sd[p + "norm.bias"], eps=1e-6)
    x = F.conv1d(x, sd[p + "proj_in.weight"], sd[p + "proj_in.bias"]).transpose(1, 2)
    for i in range(depth):
        x = _transformer_block(sd, p + "transformer_blocks.%d." % i, x, ctx)
    x = F.conv1d(x.transpose(1, 2), sd[p + "proj_out.weight"], sd[p + "proj_out.bias"])
    return x + x_in


def unet_forward(sd, x, t, latents, prefix="model."):
    """forward_no_energy (UnetInference.py:176-224) after the scene latents are known.

    x: [N,H,D]; t: [N]; latents: [N,ctx_dim] with unconditional rows already zeroed
    (UnetInference.py:190-197).  Returns f(x): [N,H,D].
    """
    t_emb = time_embedding(sd, t, prefix + "time_mlp.encoder.")
    ctx = latents[:, None, :]
    h = x.transpose(1, 2)
    skips = []
    n_levels = 4
    for i in range(n_levels):
        p = prefix + "downs.%d." % i
        h = _res_block(sd, p + "0.", h, t_emb)
        h = _res_block(sd, p + "1.", h, t_emb)
        h = _spatial_transformer(sd, p + "3.", h, ctx)
        skips.append(h)
        if i < n_levels - 1:
            h = F.conv1d(h, sd[p + "4.conv.weight"], sd[p + "4.conv.bias"], stride=2, padding=1)
    h = _res_block(sd, prefix + "mid_block1.", h, t_emb)
    h = _spatial_transformer(sd, prefix + "mid_attention.", h, ctx)
    h = _res_block(sd, prefix + "mid_block2.", h, t_emb)
    for i in range(n_levels - 1):
        p = prefix + "ups.%d." % i
        h = torch.cat([h, skips.pop()], dim=1)
        h = _res_block(sd, p + "0.", h, t_emb)
        h = _res_block(sd, p + "1.", h, t_emb)
        h = _spatial_transformer(sd, p + "3.", h, ctx)
        h = F.conv_transpose1d(h, sd[p + "4.conv.weight"], sd[p + "4.conv.bias"], stride=2, padding=1)
    h = _conv_block(sd, prefix + "final_conv.0.", h)
    h = F.conv1d(h, sd[prefix + "final_conv.1.weight"], sd[prefix + "final_conv.1.bias"])
    return h.transpose(1, 2)


def energy_grad(sd, x, t, latents, prefix="model."):
    """EnergyGradFunction (UnetInference.py:19-37): eps = d/dx 0.5*sum f(x)^2."""
    with torch.enable_grad():
        xi = x.detach().clone().requires_grad_(True)
        f = unet_forward(sd, xi, t, latents, prefix)
        e = 0.5 * (f ** 2).sum()
        (g,) = torch.autograd.grad(e, xi)
    return g.detach()


# --------------------------------------------------------------------------------------
# scene encoders (cached precompute; obstacle_encoder.py:94-152, obstacle_encoder3d.py:55-94)
# --------------------------------------------------------------------------------------


def _ln(sd, p, x):
    return F.layer_norm(x, (x.shape[-1],), sd[p + "weight"], sd[p + "bias"], eps=1e-5)


def _sincos_pe(v, div):
    """obstacle_encoder.py:73-76 / 86-89: even slots sin(x)+sin(y), odd slots cos(x)+cos(y)."""
    pe = torch.zeros(*v.shape[:-1], 2 * div.numel())
    pe[..., 0::2] = torch.sin(v[..., 0, None] * div) + torch.sin(v[..., 1, None] * div)
    pe[..., 1::2] = torch.cos(v[..., 0, None] * div) + torch.cos(v[..., 1, None] * div)
    return pe


def scene_encoder_2d(sd, pts, prefix="model.scene_encoder."):
    """ObstacleEncoderSet.forward.  pts: [S,6,64,2] -> [S,320]."""
    S, n_obs, n_pts, _ = pts.shape
    hid = 64
    div = torch.exp(torch.arange(0, hid, 2, dtype=torch.float32) * -(math.log(10000.0) / hid))
    centres = pts.mean(dim=2)
    pe_obs = _sincos_pe(centres, div)
    rel = pts - centres[:, :, None, :]
    mx = rel.abs().reshape(S, n_obs, -1).max(dim=-1, keepdim=True)[0]
    rel = rel / (mx[..., None] + 1e-8)
    pe_rel = _sincos_pe(rel, div)
    p = prefix + "point_embedding."
    emb = F.gelu(_ln(sd, p + "1.", F.linear(pts.reshape(-1, 2), sd[p + "0.weight"], sd[p + "0.bias"])))
    emb = emb.reshape(S, n_obs, n_pts, hid)
    comb = torch.cat([emb, pe_obs[:, :, None, :].expand(-1, -1, n_pts, -1), pe_rel], dim=-1)
    p = prefix + "combined_encoder."
    comb = F.gelu(_ln(sd, p + "1.", F.linear(comb, sd[p + "0.weight"], sd[p + "0.bias"])))
    tok = comb.reshape(S, n_obs * n_pts, hid)
    outs = []
    for si in range(3):
        z = tok
        for bi in range(3):
            p = prefix + "set_transformers.%d.%d." % (si, bi)
            a = _ln(sd, p + "norm1.", z)
            qkv = F.linear(a, sd[p + "attn.qkv.weight"]).reshape(S, -1, 3, 4, hid // 4).permute(2, 0, 3, 1, 4)
            q, k, v = qkv[0], qkv[1], qkv[2]
            att = (torch.matmul(q, k.transpose(-2, -1)) * ((hid // 4) ** -0.5)).softmax(dim=-1)
            o = torch.matmul(att, v).transpose(1, 2).reshape(S, -1, hid)
            z = z + F.linear(o, sd[p + "attn.proj.weight"], sd[p + "attn.proj.bias"])
            m = F.linear(_ln(sd, p + "norm2.", z), sd[p + "mlp.0.weight"], sd[p + "mlp.0.bias"])
            z = z + F.linear(F.gelu(m), sd[p + "mlp.3.weight"], sd[p + "mlp.3.bias"])
        p = prefix + "poolings.%d." % si
        pooled = F.linear(z.mean(dim=1), sd[p + "0.weight"], sd[p + "0.bias"])
        outs.append(F.linear(F.gelu(pooled), sd[p + "2.weight"], sd[p + "2.bias"]))
    return torch.cat(outs, dim=-1)


def scene_encoder_3d(sd, pts, prefix="model.scene_encoder."):
    """ObstacleEncoder.forward (eval mode: BatchNorm running stats, no dropout).
    pts: [S,n_obs,n_pts,3] -> [S,256]."""
    S, n_obs, n_pts, _ = pts.shape
    E = 256
    x = pts.reshape(-1, n_pts, 3).transpose(1, 2)
    p = prefix + "point_processor."
    for conv, bn in (("conv1", "bn1"), ("conv2", "bn2")):
        x = F.conv1d(x, sd[p + conv + ".weight"], sd[p + conv + ".bias"])
        x = F.batch_norm(x, sd[p + bn + ".running_mean"], sd[p + bn + ".running_var"],
                         sd[p + bn + ".weight"], sd[p + bn + ".bias"], training=False, eps=1e-5)
        x = F.selu(x)
    x = x.max(dim=2)[0].reshape(S, n_obs, E)
    heads, d = 4, E // 4
    for bi in range(2):
        p = prefix + "set_transformer_blocks.%d." % bi
        a = _ln(sd, p + "norm1.", x)
        qkv = F.linear(a, sd[p + "mha.in_proj_weight"], sd[p + "mha.in_proj_bias"])
        q, k, v = [t.reshape(S, n_obs, heads, d).transpose(1, 2) for t in qkv.chunk(3, dim=-1)]
        att = (torch.matmul(q, k.transpose(-2, -1)) * (d ** -0.5)).softmax(dim=-1)
        o = torch.matmul(att, v).transpose(1, 2).reshape(S, n_obs, E)
        x = x + F.linear(o, sd[p + "mha.out_proj.weight"], sd[p + "mha.out_proj.bias"])
        m = F.selu(F.linear(_ln(sd, p + "norm2.", x), sd[p + "ffn.0.weight"], sd[p + "ffn.0.bias"]))
        x = x + F.linear(m, sd[p + "ffn.3.weight"], sd[p + "ffn.3.bias"])
    feat = F.linear(x, sd[prefix + "output_proj.weight"], sd[prefix + "output_proj.bias"])
    g = feat.max(dim=1)[0]
    p = prefix + "global_pooling."
    return F.linear(F.selu(F.linear(g, sd[p + "0.weight"], sd[p + "0.bias"])), sd[p + "2.weight"], sd[p + "2.bias"])


# --------------------------------------------------------------------------------------
# score combine + sampler steps
# --------------------------------------------------------------------------------------


def inpaint(x, hard_conds):
    """apply_hard_conditioning (sample_functions.py:5-10).  hard_conds: {t: [D] or [B,D]}."""
    for t, val in hard_conds.items():
        x[:, t, :] = val
    return x


def eps_combined(sd, x, t_int, scene_latents, weights, prefix="model."):
    """CFG / compositional score sum.

    scene_latents: [G,ctx] conditional latents (G=1: CFG; G=2: reference compose; G=6: config 2).
    weights: list of G guidance weights w_g.  Returns eps_u + sum_g w_g (eps_g - eps_u), i.e.
    (1+w) eps_c - w eps_u for G=1 (diffusion_model_static.py:161-165, _dynamic.py:157-160,
    _3d.py:150) and the conjunction form at diffusion_model_static.py:213-216 / _3d.py:165-174.
    """
    B = x.shape[0]
    G = scene_latents.shape[0]
    lat = torch.cat([scene_latents, torch.zeros(1, scene_latents.shape[1])], dim=0)  # uncond row = 0
    xr = x[:, None].expand(B, G + 1, *x.shape[1:]).reshape(B * (G + 1), *x.shape[1:])
    lr = lat[None].expand(B, G + 1, -1).reshape(B * (G + 1), -1)
    t = torch.full((B * (G + 1),), int(t_int), dtype=torch.long)
    e = energy_grad(sd, xr, t, lr, prefix).reshape(B, G + 1, *x.shape[1:])
    eu = e[:, G]
    if G == 1:
        w = weights[0]
        return (1 + w) * e[:, 0] - w * eu
    out = eu.clone()
    for g in range(G):
        out = out + weights[g] * (e[:, g] - eu)
    return out


def x0_from_eps(sch, x, t, eps):
    """predict_start_from_noise + clamp (diffusion_model_static.py:109-118,167-169)."""
    x0 = sch["sqrt_recip_alphas_cumprod"][t] * x - sch["sqrt_recipm1_alphas_cumprod"][t] * eps
    return x0.clamp(-1.0, 1.0)


def ddpm_step(sch, x, t, eps, noise, noise_scale=1.0, mean_fn=None):
    """ddpm_sample_fn (sample_functions.py:19-48) after eps is known; t<0 is treated as 0.
    mean_fn: applied to the posterior mean before the noise (the APF branch of p_mean_variance,
    diffusion_model_static.py:172-181)."""
    t = max(int(t), 0)
    x0 = x0_from_eps(sch, x, t, eps)
    mean = sch["posterior_mean_coef1"][t] * x0 + sch["posterior_mean_coef2"][t] * x
    if mean_fn is not None:
        mean = mean_fn(mean)
    std = torch.exp(0.5 * sch["posterior_log_variance_clipped"][t])
    if t == 0:
        noise = torch.zeros_like(noise)
    return mean + std * noise * noise_scale


def ddim_step(sch, x, t, t_prev, x0):
    """ddim_p_sample with eta=0, use_clipped_model_output (diffusion_model_static.py:321-332)."""
    a_t = sch["alphas_cumprod"][int(t)]
    a_prev = sch["alphas_cumprod"][int(t_prev)] if t_prev >= 0 else torch.tensor(1.0)
    e = (x - a_t ** 0.5 * x0) / (1 - a_t) ** 0.5
    return a_prev ** 0.5 * x0 + (1 - a_prev) ** 0.5 * e


def q_sample(sch, x0, t, noise):
    """diffusion_model_dynamic.py:671-680."""
    return sch["sqrt_alphas_cumprod"][int(t)] * x0 + sch["sqrt_one_minus_alphas_cumprod"][int(t)] * noise


# --------------------------------------------------------------------------------------
# nearest point / APF  (APFhelper.py, APFhelper_dynamic.py)
# --------------------------------------------------------------------------------------


def nn_query(cloud, q, upper_bound):
    """k=1 nearest neighbour with cKDTree.query(distance_upper_bound=) semantics
    (APFhelper.py:17-35): float64 arithmetic, miss -> (inf, P).  cloud [P,d], q [M,d] arrays."""
    cloud = np.asarray(cloud, dtype=np.float64)
    q = np.asarray(q, dtype=np.float64)
    P = cloud.shape[0]
    dist = np.full(q.shape[0], np.inf)
    idx = np.full(q.shape[0], P, dtype=np.int64)
    step = max(1, (1 << 24) // max(P, 1))
    for s in range(0, q.shape[0], step):
        d2 = ((q[s:s + step, None, :] - cloud[None, :, :]) ** 2).sum(-1)
        j = d2.argmin(axis=1)
        d = np.sqrt(d2[np.arange(d2.shape[0]), j])
        hit = d < upper_bound
        dist[s:s + step][hit] = d[hit]
        idx[s:s + step][hit] = j[hit]
    return dist, idx


def apf_static(x, cloud, thr=0.07, strength=0.1, window=7, pos_dims=2):
    """avoidance (APFhelper.py:37-104).  x: [B,H,D] tensor, cloud: [P,d] tensor -> new [B,H,D]."""
    B, H, _ = x.shape
    d = pos_dims
    dist, idx = nn_query(cloud.numpy(), x[..., :d].reshape(-1, d).numpy(), thr)
    if dist.min() > thr:
        return x
    dist = dist.reshape(B, H)
    idx = idx.reshape(B, H)
    ks = torch.arange(-window, window + 1)
    wts = torch.exp(-0.5 * ks.float() ** 2 / (window / 2) ** 2)
    force = torch.zeros(B, H, d)
    hit = (dist < thr) & (idx < cloud.shape[0])
    for b, t in zip(*np.nonzero(hit)):
        direction = x[b, t, :d] - cloud[idx[b, t]]
        direction = direction / (direction.norm() + 1e-8)
        mag = strength * math.exp(-dist[b, t] / thr)
        for k, w in zip(ks.tolist(), wts):
            if 0 <= t + k < H:
                force[b, t + k] += (mag * direction.double() * w.double()).float()
    out = x.clone()
    out[..., :d] += force
    return out


def apf_dynamic(traj, pts, thr_query, thr_force, strength, window=None, is_dynamic=False,
                affected=None, goal=None):
    """Per-trajectory avoidance (APFhelper_dynamic.py:107-142), in place on a copy.

    traj [H,D]; pts [P,2] numpy (static cloud or pursuer disc); thr_query = the KD-tree upper
    bound of that field; thr_force = ``obstacle_field.distance_threshold`` (the static one is used
    for both fields, :139).  Static mode pushes [c-w, min(H-1, c+w)) around the closest index c;
    dynamic mode pushes [0, affected)."""
    traj = traj.clone()
    H = traj.shape[0]
    if is_dynamic:
        n = H if affected is None else min(affected, H)
        dist, idx = nn_query(pts, traj[:n, :2].numpy(), thr_query)
        lo, hi = 0, n
    else:
        dist, idx = nn_query(pts, traj[:, :2].numpy(), thr_query)
        c = int(np.argmin(dist))
        lo, hi = max(0, c - window), min(H - 1, c + window)
    for i in range(lo, hi):
        if idx[i] < len(pts):
            away = traj[i, :2] - torch.from_numpy(np.asarray(pts[idx[i]])).to(traj.dtype)
            away = away / (away.norm() + 1e-8)
            if goal is not None:
                tog = goal[:2] - traj[i, :2]
                tog = tog / (tog.norm() + 1e-8)
                direction = 0.9 * away + 0.1 * tog
                direction = direction / (direction.norm() + 1e-8)
            else:
                direction = away
            traj[i, :2] += (strength * math.exp(-dist[i] / thr_force) * direction.double()).float()
    return traj


# --------------------------------------------------------------------------------------
# cost + selection (cost.py)
# --------------------------------------------------------------------------------------


def collision_mask(x, cloud, thr, pos_dims=2):
    """compute_collision_with_pointcloud (cost.py:25-54): any ||p - q|| < thr.  cloud [P,d]."""
    out = torch.zeros(x.shape[0], dtype=torch.bool)
    q = x[..., :pos_dims]
    for b in range(x.shape[0]):
        dmat = torch.norm(q[b][:, None, :] - cloud[None, :, :], dim=-1)
        out[b] = bool((dmat < thr).any())
    return out


def path_length(x, pos_dims=2):
    """cost.py:3-7."""
    return torch.linalg.norm(torch.diff(x[..., :pos_dims], dim=-2), dim=-1).sum(-1)


def smoothness(x, pos_dims=2):
    """cost.py:19-24 (velocity channels follow the position channels)."""
    return torch.linalg.norm(torch.diff(x[..., pos_dims:], dim=-2), dim=-1).sum(-1)


def trajectory_costs(x, cloud, thr, w_smooth=0.1, w_len=0.9, pos_dims=2):
    """compute_trajectory_costs (cost.py:56-88).  Returns the reference 5-tuple."""
    free = ~collision_mask(x, cloud, thr, pos_dims)
    if not free.any():
        return None, None, None, free, None
    xs = x[free]
    pl, sm = path_length(xs, pos_dims), smoothness(xs, pos_dims)
    pl = (pl - pl.min()) / (pl.max() - pl.min())
    sm = (sm - sm.min()) / (sm.max() - sm.min())
    cost = w_smooth * sm + w_len * pl
    j = torch.argmin(cost)
    return xs[j], cost[j], cost, free, j


def smooth_segment(s1, s2, num_steps, dt=0.1, max_vel=0.8):
    """DynamicGaussianDiffusionModel.sm (diffusion_model_dynamic.py:192-214)."""
    dp = s2[:, :2] - s1[:, :2]
    dist = dp.norm(dim=1, keepdim=True)
    direction = torch.where(dist > 1e-6, dp / dist, torch.zeros_like(dp))
    want = dp / (num_steps * dt)
    v = torch.where(want.norm(dim=1, keepdim=True) > max_vel, direction * max_vel, want)
    tt = (torch.arange(1, num_steps + 1).float() * dt).view(1, num_steps, 1)
    pos = s1[:, None, :2] + tt * v[:, None, :]
    return torch.cat([pos, v[:, None, :].expand(-1, num_steps, -1)], dim=-1)


# --------------------------------------------------------------------------------------
# whole loops, with an explicit noise list (noise[0] = x_T, noise[1+k] = step-k draw)
# --------------------------------------------------------------------------------------


def sample_ddpm(sd, sch, T, scene_latents, weights, hard_conds, noise, noise_scale=0.5,
                extra_steps=0, prefix="model.", apf_cloud=None, apf_after=20, apf_thr=0.07, apf_strength=0.1,
                apf_window=5):
    """p_sample_loop + ddpm_sample_fn (diffusion_model_static.py:231-256, _3d.py:184-218).
    apf_cloud [P, 2]: the static model's `use_apf` DDPM branch — for forward_t > 20 the posterior mean goes through
    `avoidance` (threshold 0.07, strength 0.1, window 5) before the noise is added (diffusion_model_static.py:172-181;
    CFG path only, p_mean_variance_compose has no such branch)."""
    x = inpaint(noise[0].clone(), hard_conds)
    chain = [x.clone()]
    k = 1
    for fwd, i in enumerate(reversed(range(-extra_steps, T))):
        t = max(i, 0)
        eps = eps_combined(sd, x, t, scene_latents, weights, prefix)
        push = None
        if apf_cloud is not None and fwd > apf_after:
            push = lambda m: apf_static(m, apf_cloud, apf_thr, apf_strength, apf_window)  # noqa: E731
        x = ddpm_step(sch, x, t, eps, noise[k], noise_scale, mean_fn=push)
        k += 1
        x = inpaint(x, hard_conds)
        chain.append(x.clone())
    return torch.stack(chain, 0)


def sample_ddim(sd, sch, T, n_steps, scene_latents, weights, hard_conds, noise, apf_cloud=None,
                apf_from=2, apf_iters=3, apf_thr=0.07, apf_strength=0.1, apf_window=7, prefix="model."):
    """ddim_p_sample_loop (diffusion_model_static.py:347-384) incl. the APF hook (:298-319)."""
    x = inpaint(noise[0].clone(), hard_conds)
    chain = [x.clone()]
    for fwd, t in enumerate(ddim_timesteps(T, n_steps)):
        eps = eps_combined(sd, x, int(t), scene_latents, weights, prefix)
        x0 = x0_from_eps(sch, x, int(t), eps)
        if apf_cloud is not None and fwd >= apf_from:
            for _ in range(apf_iters):
                x0 = apf_static(x0, apf_cloud, apf_thr, apf_strength, apf_window)
                x0 = inpaint(x0, hard_conds)
        x = ddim_step(sch, x, int(t), int(t) - T // n_steps, x0)
        x = inpaint(x, hard_conds)
        chain.append(x.clone())
    return torch.stack(chain, 0)


# --------------------------------------------------------------------------------------
# dynamic model: stacked-CFG row quirk, stage I, one replan tick (diffusion_model_dynamic.py)
# --------------------------------------------------------------------------------------


def eps_dynamic(sd, x, t_int, scene_latent, w=2.5, prefix="model."):
    """p_mean_variance of the dynamic model (diffusion_model_dynamic.py:152-160): the batch is stacked twice and the
    network zeroes the latent of every odd row of the stack (UnetInference.py:192-195)."""
    B = x.shape[0]
    x2 = x.repeat(2, 1, 1)
    lat = scene_latent.reshape(1, -1).repeat(2 * B, 1).clone()
    lat[1::2] = 0
    t = torch.full((2 * B,), int(t_int), dtype=torch.long)
    out = energy_grad(sd, x2, t, lat, prefix)
    return (1 + w) * out[:B] - w * out[B:]


def dynamic_stage1(sd, sch, T, scene_latent, hard_conds, x_T, n_high=10):
    """Stage I (diffusion_model_dynamic.py:537-546): n_high DDIM steps without APF."""
    x = inpaint(x_T.clone(), hard_conds)
    for t in ddim_timesteps(T, n_high):
        eps = eps_dynamic(sd, x, int(t), scene_latent)
        x0 = x0_from_eps(sch, x, int(t), eps)
        x = inpaint(ddim_step(sch, x, int(t), int(t) - T // n_high, x0), hard_conds)
    return x


def _pin_history(x, history, goal_row, zero_start_velocity):
    for i, h in enumerate(history):
        x[:, i] = h.reshape(1, -1)
    x[:, -1] = goal_row
    if zero_start_velocity:
        x[:, 0, 2:] = 0.0
    return x


def dynamic_tick_batch(sd, sch, T, scene_latent, x_clean, history, stepp, k, noise, static_pts, disc_pts, center,
                       n_high=10, n_low=5, thr_static=0.2, thr_pred=0.5, strength=0.15, w_static=8, w_pred=5):
    """One stage-II replan up to the selection (diffusion_model_dynamic.py:553-591).  history: list of [1, D]
    executed states; static_pts / disc_pts: numpy clouds of the ObstacleField; center: pursuer centre (2,)."""
    H = x_clean.shape[0]
    ts = ddim_timesteps(T, n_high)
    low = ts[-n_low:]
    goal_row = x_clean[-1].clone()
    x = q_sample(sch, x_clean[None].repeat(noise.shape[0], 1, 1), int(low[0]), noise)
    x[:, 0, 2:] = 0
    x = _pin_history(x, history, goal_row, zero_start_velocity=False)
    for t in low:
        t = int(t)
        if t == 0:
            x[:, stepp + 1:stepp + 4] = smooth_segment(x[:, stepp], x[:, stepp + 3], 3)
        eps = eps_dynamic(sd, x, t, scene_latent)
        x0 = x0_from_eps(sch, x, t, eps)
        if t == 0:
            start = x[:, k].clone()
            dist = (start[:, :2] - torch.as_tensor(center, dtype=x.dtype)[None]).norm(dim=1)
            for b in range(x.shape[0]):
                x0[b] = apf_dynamic(x0[b], static_pts, thr_static, thr_static, strength, window=w_static)
                if dist[b] < thr_pred:
                    x0[b] = apf_dynamic(x0[b], disc_pts, thr_pred, thr_static, strength, window=w_pred, is_dynamic=True,
                                        affected=H, goal=x[0, -1])
            x0[:, -1] = x[:, -1]
        x = ddim_step(sch, x, t, t - T // n_high, x0)
        x = _pin_history(x, history, goal_row, zero_start_velocity=True)
    x[:, stepp + 1:stepp + 3] = smooth_segment(x[:, stepp], x[:, stepp + 2], 2)
    return x


# --------------------------------------------------------------------------------------
# evaluation metrics (scripts/inference/core/metrics.py)
# --------------------------------------------------------------------------------------

def metric_collision_intensity(x, centers, sizes):
    """Metrics.compute_collision_intensity (metrics.py:48-82): fraction of waypoints inside any inclusive box."""
    centers = torch.as_tensor(centers, dtype=torch.float32).reshape(-1, 2)
    sizes = torch.as_tensor(sizes, dtype=torch.float32)
    if sizes.dim() == 1:
        sizes = sizes[:, None].repeat(1, 2)
    lo, hi = centers - sizes / 2, centers + sizes / 2
    p = x[:, :, None, :2]
    inside = ((p >= lo) & (p <= hi)).all(dim=-1)
    return inside.any(dim=-1).float().mean(dim=1)


def metric_variance_waypoints(x, eps=1e-8):
    """Metrics.compute_variance_waypoints (metrics.py:9-19), restated with exact fp64 pair distances: per horizon step the
    unbiased variance over ALL n*n entries of triu(cdist, 1) + eps (zeros included), summed over the horizon."""
    p = x[..., :2].double().permute(1, 0, 2)                       # [H, n, 2]
    d = (p[:, :, None, :] - p[:, None, :, :]).norm(dim=-1)
    d = torch.triu(d, diagonal=1).reshape(d.shape[0], -1) + eps
    return d.var(dim=1).sum().float()


def metric_success(x, ci, thr=0.01):
    """Metrics.trajectory_success_and_metrics (metrics.py:84-126), scalar fields only."""
    ok = ci <= thr
    free = x[ok]
    out = {"success": int(ok.any()), "collision_intensity": ci.mean().item() * 100, "n_free_trajectories": len(free),
           "path_length": None, "path_length_std": None, "waypoint_variance": None}
    if len(free) > 0:
        pl = path_length(free, 2)
        out["path_length"], out["path_length_std"] = pl.mean().item(), pl.std().item()
        out["waypoint_variance"] = 0.0 if len(free) == 1 else metric_variance_waypoints(free).item()
    return out


# --------------------------------------------------------------------------------------
# caller-side state of the dynamic planner, with numpy's random draws as INPUTS
# --------------------------------------------------------------------------------------
def pursuer_step(t, center, robot, velocity, pursuit_strength=0.8, random_strength=0.2, dt=0.1):
    """MultiSphereFieldDynamics.update_centers (deps/torch_robotics/torch_robotics/environments/primitives.py:100-107) with
    dynamics_fn of scripts/inference/core/utils.py:92-135: `to_numpy` hands the law float32 arrays; the unit vector to the
    robot stays float32, the wobble (sin 2πt, cos 2πt) and the position update are float64, the result is stored back as
    float32.  center / robot / velocity: 2 values each.  Returns the new centre, float32 [2]."""
    c = np.asarray(center, dtype=np.float32).reshape(-1)[:2]
    r = np.asarray(robot, dtype=np.float32).reshape(-1)[:2]
    v = np.asarray(velocity, dtype=np.float32).reshape(-1)[:2]
    d = r - c
    dist = np.sqrt(d[0] ** 2 + d[1] ** 2)
    if dist > 0:
        d = d / dist
    wob = np.array([np.sin(2 * np.pi * t), np.cos(2 * np.pi * t)])
    head = np.float32(pursuit_strength) * d + random_strength * wob
    new = c + head * v * dt
    return np.clip(new, -1, 1).astype(np.float32)


def disc_points(center, radius, num_points, draws, surface_ratio=0.9):
    """generate_sphere_points (mpd/models/diffusion_models/APFhelper_dynamic.py:18-39); draws = the n_in U(0,1) radius draws
    followed by the n_in U(0,2π) angle draws the reference takes from np.random (:28-29).  Returns float64 [num_points, 2]."""
    n_rim = int(num_points * surface_ratio)
    n_in = num_points - n_rim
    th = np.pi * (3 - np.sqrt(5)) * np.arange(n_rim)
    r = np.full(n_rim, float(radius))
    if n_in > 0:
        draws = np.asarray(draws, dtype=np.float64)
        r = np.concatenate((r, float(radius) * np.sqrt(draws[:n_in])))
        th = np.concatenate((th, draws[n_in:2 * n_in]))
    return np.column_stack((r * np.cos(th) + center[0], r * np.sin(th) + center[1]))


def box_points(center, size, num_points, n_edge, u_edge, u_in):
    """generate_box_points (APFhelper_dynamic.py:41-68) for one box with float32 centre / size (what the dynamic loop passes,
    diffusion_model_dynamic.py:516-517); n_edge = its randint draw (:49), u_edge = rand(n_edge) (:55), u_in = rand(n_in, 2)
    (:64).  The perimeter is walked over the reference's 8-entry length table [w, w, h, h, w, w, h, h]."""
    c = np.asarray(center, dtype=np.float32)
    s = np.asarray(size, dtype=np.float32)
    w, h = s[0], s[1]
    left, right, top, bottom = c[0] - w / 2, c[0] + w / 2, c[1] + h / 2, c[1] - h / 2
    corners = np.array([[left, top], [right, top], [right, bottom], [left, bottom]])
    lens = np.array([w, h, w, h]).repeat(2)
    along = np.asarray(u_edge, dtype=np.float64)[:n_edge] * lens.sum()
    cum = np.cumsum(lens)
    seg = np.searchsorted(cum, along)
    frac = (along - np.concatenate(([0], cum[:-1]))[seg]) / lens[seg]
    a, b = corners[seg % 4], corners[(seg + 1) % 4]
    edge = a + frac[:, None] * (b - a)
    inner = np.array(np.asarray(u_in, dtype=np.float64)[:num_points - n_edge], copy=True)
    inner[:, 0] = inner[:, 0] * w + left
    inner[:, 1] = inner[:, 1] * h + bottom
    return np.concatenate([edge, inner], axis=0)
