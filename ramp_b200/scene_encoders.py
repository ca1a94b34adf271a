"""Scene encoders (point cloud -> conditioning latent).

Evaluated once per distinct scene and cached (reference: UnetInference.py:146-156), or every tick when the
conditioning cloud moves.  The modules below only HOLD the parameters, under the reference's names
and shapes so its checkpoints load unchanged:
  2-D: mpd/models/diffusion_models/obstacle_encoder.py:94-152  (ObstacleEncoderSet) -> csrc/scene_enc.cu through
       ramp_scene2d_encode (8 launches, FP32); CUDA tensors only, no CPU path
  3-D: mpd/models/diffusion_models/obstacle_encoder3d.py:55-94 (ObstacleEncoder) -> PyTorch module on the model's
       device (0.03 GF per scene)
"""
import ctypes
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from ._lib import RampError, check, lib, ptr, stream_ptr


class _SetAttention(nn.Module):
    """Parameters of the 4-head set attention: ``qkv`` (no bias) / ``proj`` (obstacle_encoder.py:6-32)."""

    def __init__(self, dim, heads=4):
        super().__init__()
        self.heads = heads
        self.qkv = nn.Linear(dim, 3 * dim, bias=False)
        self.proj = nn.Linear(dim, dim)


class _SetBlock(nn.Module):
    """Parameters of a pre-norm transformer block; ``mlp`` indices 0 and 3 hold the linears (obstacle_encoder.py:34-50)."""

    def __init__(self, dim):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn = _SetAttention(dim)
        self.norm2 = nn.LayerNorm(dim)
        self.mlp = nn.Sequential(nn.Linear(dim, 4 * dim), nn.GELU(), nn.Identity(), nn.Linear(4 * dim, dim))


class _CentrePE(nn.Module):
    """``div_term`` buffer of the sinusoidal code of obstacle centres / max-normalised offsets (obstacle_encoder.py:52-92)."""

    def __init__(self, d_model):
        super().__init__()
        self.register_buffer("div_term", torch.exp(torch.arange(0, d_model, 2) * -(math.log(10000.0) / d_model)))


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

    def _flat_params(self):
        """The state-dict tensors concatenated in the C ABI's order (rebuilt when a parameter changed or moved)."""
        L = lib()
        L.ramp_scene2d_param_name.restype = ctypes.c_char_p
        L.ramp_scene2d_param_numel.restype = ctypes.c_size_t
        sd = dict(self.state_dict())
        key = tuple((t.data_ptr(), t._version) for t in sd.values())
        if getattr(self, "_flat_key", None) != key:
            parts = []
            for i in range(L.ramp_scene2d_param_count()):
                name = L.ramp_scene2d_param_name(i).decode()
                t = sd[name]
                if t.numel() != L.ramp_scene2d_param_numel(i):
                    raise RampError("scene encoder parameter %s has %d elements, expected %d"
                                    % (name, t.numel(), L.ramp_scene2d_param_numel(i)))
                parts.append(t.detach().reshape(-1).to(torch.float32))
            self._flat = torch.cat(parts).contiguous()
            self._flat_key = key
        return self._flat

    def forward(self, pts):
        if not pts.is_cuda:
            raise RampError("ObstacleEncoderSet runs on CUDA tensors only (csrc/scene_enc.cu); there is no CPU path")
        if tuple(pts.shape[1:]) != (self.num_obstacles, self.num_points, 2) or self.num_obstacles != 6 \
                or self.num_points != 64:
            raise RampError("ObstacleEncoderSet is hard-wired to clouds [S, 6, 64, 2] (obstacle_encoder.py:95,132-134)")
        L = lib()
        L.ramp_scene2d_workspace_floats.restype = ctypes.c_size_t
        S = pts.shape[0]
        pts = pts.detach().to(torch.float32).contiguous()
        flat = self._flat_params()
        ws = torch.empty(max(1, L.ramp_scene2d_workspace_floats(S)), device=pts.device)
        out = torch.empty(S, 320, device=pts.device)
        check(L.ramp_scene2d_encode(ptr(flat), ptr(pts), S, ptr(ws), ptr(out), stream_ptr()), "ramp_scene2d_encode")
        return out


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
