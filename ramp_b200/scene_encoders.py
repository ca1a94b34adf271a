"""Scene encoders (point cloud -> conditioning latent).

Cached precompute, evaluated once per distinct scene and never inside the diffusion loop
(reference: UnetInference.py:146-156 caches the encoder output).  They stay as PyTorch modules
on the model's device; the per-step hot path is CUDA only.
Parameter names and shapes match the reference so its checkpoints load unchanged:
  2-D: mpd/models/diffusion_models/obstacle_encoder.py:94-152  (ObstacleEncoderSet)
  3-D: mpd/models/diffusion_models/obstacle_encoder3d.py:55-94 (ObstacleEncoder)
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class _SetAttention(nn.Module):
    """4-head attention over set tokens; names ``qkv`` (no bias) / ``proj`` (obstacle_encoder.py:6-32)."""

    def __init__(self, dim, heads=4):
        super().__init__()
        self.heads = heads
        self.qkv = nn.Linear(dim, 3 * dim, bias=False)
        self.proj = nn.Linear(dim, dim)

    def forward(self, x):
        S, n, c = x.shape
        q, k, v = self.qkv(x).view(S, n, 3, self.heads, c // self.heads).permute(2, 0, 3, 1, 4)
        o = F.scaled_dot_product_attention(q, k, v)  # scale = head_dim ** -0.5
        return self.proj(o.transpose(1, 2).reshape(S, n, c))


class _SetBlock(nn.Module):
    """Pre-norm transformer block; ``mlp`` indices 0 and 3 hold the linears (obstacle_encoder.py:34-50)."""

    def __init__(self, dim):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn = _SetAttention(dim)
        self.norm2 = nn.LayerNorm(dim)
        self.mlp = nn.Sequential(nn.Linear(dim, 4 * dim), nn.GELU(), nn.Identity(), nn.Linear(4 * dim, dim))

    def forward(self, x):
        x = x + self.attn(self.norm1(x))
        return x + self.mlp(self.norm2(x))


class _CentrePE(nn.Module):
    """Sinusoidal code of obstacle centres and of max-normalised offsets (obstacle_encoder.py:52-92)."""

    def __init__(self, d_model):
        super().__init__()
        self.register_buffer("div_term", torch.exp(torch.arange(0, d_model, 2) * -(math.log(10000.0) / d_model)))

    def _code(self, v):
        s = torch.sin(v[..., 0:1] * self.div_term) + torch.sin(v[..., 1:2] * self.div_term)
        c = torch.cos(v[..., 0:1] * self.div_term) + torch.cos(v[..., 1:2] * self.div_term)
        return torch.stack([s, c], dim=-1).flatten(-2)  # even slots = sin sum, odd slots = cos sum

    def forward(self, pts):
        centres = pts.mean(dim=2)
        rel = pts - centres.unsqueeze(2)
        peak = rel.abs().flatten(2).amax(dim=-1)
        rel = rel / (peak[..., None, None] + 1e-8)
        return self._code(centres), self._code(rel)


class ObstacleEncoderSet(nn.Module):
    """[S, 6, 64, 2] -> [S, 320] (three set-transformer stacks pooled to 64/96/160)."""

    def __init__(self, input_dim=2, hidden_dim=64, output_dims=(64, 96, 160), num_obstacles=6, num_points=64,
                 num_blocks=3):
        super().__init__()
        self.num_obstacles, self.num_points = num_obstacles, num_points
        self.pos_encoder = _CentrePE(hidden_dim)
        self.point_embedding = nn.Sequential(nn.Linear(input_dim, hidden_dim), nn.LayerNorm(hidden_dim), nn.GELU())
        self.combined_encoder = nn.Sequential(nn.Linear(3 * hidden_dim, hidden_dim), nn.LayerNorm(hidden_dim),
                                              nn.GELU())
        self.set_transformers = nn.ModuleList(
            [nn.Sequential(*[_SetBlock(hidden_dim) for _ in range(num_blocks)]) for _ in output_dims])
        self.poolings = nn.ModuleList(
            [nn.Sequential(nn.Linear(hidden_dim, d), nn.GELU(), nn.Linear(d, d)) for d in output_dims])

    def forward(self, pts):
        S = pts.shape[0]
        pe_obs, pe_rel = self.pos_encoder(pts)
        emb = self.point_embedding(pts)
        tok = torch.cat([emb, pe_obs.unsqueeze(2).expand(-1, -1, self.num_points, -1), pe_rel], dim=-1)
        tok = self.combined_encoder(tok).reshape(S, self.num_obstacles * self.num_points, -1)
        return torch.cat([pool(stack(tok).mean(dim=1)) for stack, pool in zip(self.set_transformers, self.poolings)],
                         dim=-1)


class _PointNet(nn.Module):
    """Shared 1x1 convs + eval-mode BatchNorm + SELU + max over points (obstacle_encoder3d.py:5-21)."""

    def __init__(self, input_dim, output_dim):
        super().__init__()
        self.conv1 = nn.Conv1d(input_dim, 64, 1)
        self.conv2 = nn.Conv1d(64, output_dim, 1)
        self.bn1 = nn.BatchNorm1d(64)
        self.bn2 = nn.BatchNorm1d(output_dim)

    def forward(self, x):  # [S*n_obs, n_pts, 3]
        x = x.transpose(1, 2)
        x = F.selu(self.bn1(self.conv1(x)))
        x = F.selu(self.bn2(self.conv2(x)))
        return x.amax(dim=2)


class _ObstacleBlock(nn.Module):
    """nn.MultiheadAttention block over obstacles (obstacle_encoder3d.py:23-53); ``ffn`` indices 0/3."""

    def __init__(self, dim, heads=4):
        super().__init__()
        self.mha = nn.MultiheadAttention(dim, heads, dropout=0.1)
        self.ffn = nn.Sequential(nn.Linear(dim, 2 * dim), nn.SELU(), nn.Identity(), nn.Linear(2 * dim, dim))
        self.norm1 = nn.LayerNorm(dim)
        self.norm2 = nn.LayerNorm(dim)

    def forward(self, x):  # [S, n_obs, dim]
        a = self.norm1(x).transpose(0, 1)
        x = x + self.mha(a, a, a, need_weights=False)[0].transpose(0, 1)
        return x + self.ffn(self.norm2(x))


class ObstacleEncoder(nn.Module):
    """[S, n_obs, n_pts, 3] -> [S, 256]."""

    def __init__(self, embedding_dim=256, point_dim=3, num_layers=2):
        super().__init__()
        self.embedding_dim = embedding_dim
        self.point_processor = _PointNet(point_dim, embedding_dim)
        self.set_transformer_blocks = nn.ModuleList([_ObstacleBlock(embedding_dim) for _ in range(num_layers)])
        self.output_proj = nn.Linear(embedding_dim, embedding_dim)
        self.global_pooling = nn.Sequential(nn.Linear(embedding_dim, embedding_dim), nn.SELU(),
                                            nn.Linear(embedding_dim, embedding_dim))

    def forward(self, pts):
        S, n_obs, n_pts, dim = pts.shape
        x = self.point_processor(pts.reshape(-1, n_pts, dim)).view(S, n_obs, self.embedding_dim)
        for blk in self.set_transformer_blocks:
            x = blk(x)
        return self.global_pooling(self.output_proj(x).amax(dim=1))
