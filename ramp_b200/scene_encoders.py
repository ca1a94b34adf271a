"""Scene encoders (point cloud -> conditioning latent).

Evaluated once per distinct scene and cached (reference: UnetInference.py:146-156), or every tick when the
conditioning cloud moves.  The modules below only HOLD the parameters, under the reference's names
and shapes so its checkpoints load unchanged:
  2-D: mpd/models/diffusion_models/obstacle_encoder.py:94-152  (ObstacleEncoderSet) -> csrc/scene_enc.cu through
       ramp_scene2d_encode (8 launches, FP32); CUDA tensors only, no CPU path
  3-D: mpd/models/diffusion_models/obstacle_encoder3d.py:55-94 (ObstacleEncoder) -> ramp_scene3d_encode (2 launches)
"""
import ctypes
import math

import torch
import torch.nn as nn

from ._lib import RampError, check, lib, ptr, stream_ptr


def _flat_params(module, which):
    """The module's state-dict tensors concatenated in the C ABI's order (rebuilt when a tensor changed or moved)."""
    L = lib()
    count = getattr(L, "ramp_scene%s_param_count" % which)
    name_of = getattr(L, "ramp_scene%s_param_name" % which)
    numel_of = getattr(L, "ramp_scene%s_param_numel" % which)
    name_of.restype = ctypes.c_char_p
    numel_of.restype = ctypes.c_size_t
    sd = dict(module.state_dict())
    key = tuple((t.data_ptr(), t._version) for t in sd.values())
    if getattr(module, "_flat_key", None) != key:
        parts = []
        for i in range(count()):
            name = name_of(i).decode()
            t = sd[name]
            if t.numel() != numel_of(i):
                raise RampError("scene encoder parameter %s has %d elements, expected %d" % (name, t.numel(), numel_of(i)))
            parts.append(t.detach().reshape(-1).to(torch.float32))
        module._flat = torch.cat(parts).contiguous()
        module._flat_key = key
    return module._flat


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
        flat = _flat_params(self, "2d")
        ws = torch.empty(max(1, L.ramp_scene2d_workspace_floats(S)), device=pts.device)
        out = torch.empty(S, 320, device=pts.device)
        check(L.ramp_scene2d_encode(ptr(flat), ptr(pts), S, ptr(ws), ptr(out), stream_ptr()), "ramp_scene2d_encode")
        return out


class _PointNet(nn.Module):
    """Parameters of the shared 1x1 convs + BatchNorm (obstacle_encoder3d.py:5-21)."""

    def __init__(self, input_dim, output_dim):
        super().__init__()
        self.conv1 = nn.Conv1d(input_dim, 64, 1)
        self.conv2 = nn.Conv1d(64, output_dim, 1)
        self.bn1 = nn.BatchNorm1d(64)
        self.bn2 = nn.BatchNorm1d(output_dim)


class _ObstacleBlock(nn.Module):
    """Parameters of the nn.MultiheadAttention block over obstacles (obstacle_encoder3d.py:23-53); ``ffn`` indices 0/3."""

    def __init__(self, dim, heads=4):
        super().__init__()
        self.mha = nn.MultiheadAttention(dim, heads, dropout=0.1)
        self.ffn = nn.Sequential(nn.Linear(dim, 2 * dim), nn.SELU(), nn.Identity(), nn.Linear(2 * dim, dim))
        self.norm1 = nn.LayerNorm(dim)
        self.norm2 = nn.LayerNorm(dim)


class ObstacleEncoder(nn.Module):
    """[S, n_obs, n_pts, 3] -> [S, 256] (eval-mode semantics: BatchNorm running statistics, no dropout)."""

    def __init__(self, embedding_dim=256, point_dim=3, num_layers=2):
        super().__init__()
        self.embedding_dim = embedding_dim
        self.point_processor = _PointNet(point_dim, embedding_dim)
        self.set_transformer_blocks = nn.ModuleList([_ObstacleBlock(embedding_dim) for _ in range(num_layers)])
        self.output_proj = nn.Linear(embedding_dim, embedding_dim)
        self.global_pooling = nn.Sequential(nn.Linear(embedding_dim, embedding_dim), nn.SELU(),
                                            nn.Linear(embedding_dim, embedding_dim))

    def forward(self, pts):
        if not pts.is_cuda:
            raise RampError("ObstacleEncoder runs on CUDA tensors only (csrc/scene_enc.cu); there is no CPU path")
        if pts.dim() != 4 or pts.shape[-1] != 3 or self.embedding_dim != 256 or len(self.set_transformer_blocks) != 2:
            raise RampError("ObstacleEncoder expects clouds [S, n_obs, n_pts, 3] and the reference's 256-wide, 2-block model")
        S, n_obs, n_pts, _ = pts.shape
        if not 1 <= n_obs <= 32:
            raise RampError("ObstacleEncoder kernels hold at most 32 obstacles per scene (got %d)" % n_obs)
        L = lib()
        L.ramp_scene3d_workspace_floats.restype = ctypes.c_size_t
        pts = pts.detach().to(torch.float32).contiguous()
        flat = _flat_params(self, "3d")
        ws = torch.empty(max(1, L.ramp_scene3d_workspace_floats(S, n_obs)), device=pts.device)
        out = torch.empty(S, 256, device=pts.device)
        check(L.ramp_scene3d_encode(ptr(flat), ptr(pts), S, n_obs, n_pts, ptr(ws), ptr(out), stream_ptr()),
              "ramp_scene3d_encode")
        return out
