"""Evaluation metrics of a sampled batch, on the device (drop-in for scripts/inference/core/metrics.py).

`Metrics` / `DynamicMetrics` keep the reference's method names, arguments and return types
(metrics.py:5-126, 128-171); the per-trajectory arithmetic runs in csrc/metrics_ops.cu through the C ABI.  CUDA
tensors only.  `calculate_single_episode_metrics` works on the host-side chains the dynamic planner returns (a few
dozen 2-vectors), as in the reference."""
import ctypes

import numpy as np
import torch

from ._lib import RampError, check, lib, ptr, stream_ptr


def _cuda3(t, what):
    if not torch.is_tensor(t) or t.dim() != 3:
        raise RampError("%s expects a [batch, horizon, state_dim] tensor" % what)
    if not t.is_cuda:
        raise RampError("%s runs on CUDA tensors only; there is no CPU path" % what)
    return t.detach().to(torch.float32).contiguous()


def _path_sums(x, c0, c1):
    B, H, D = x.shape
    out = torch.empty(B, device=x.device)
    check(lib().ramp_path_sums(ptr(x), B, H, D, c0, c1, ptr(out), stream_ptr()), "ramp_path_sums")
    return out


class Metrics:
    """metrics.py:5-126."""

    @staticmethod
    def compute_variance_waypoints(trajs, eps=1e-8):
        """metrics.py:9-19 — returns a 0-dim tensor (0.0 for fewer than two trajectories, where the reference's
        variance of a single entry is NaN; its caller never asks, metrics.py:117-118)."""
        x = _cuda3(trajs, "compute_variance_waypoints")
        n, H, D = x.shape
        out = torch.zeros(1, device=x.device)
        if n >= 2:
            L = lib()
            L.ramp_waypoint_variance_workspace_bytes.restype = ctypes.c_size_t
            ws = torch.empty(L.ramp_waypoint_variance_workspace_bytes(n, H) // 8, dtype=torch.float64, device=x.device)
            check(L.ramp_waypoint_variance(ptr(x), n, H, D, ptr(ws), ptr(out), stream_ptr()), "ramp_waypoint_variance")
        return out[0]

    @staticmethod
    def compute_smoothness(trajs, trajs_vel=None):
        """metrics.py:21-31 — [batch] sums of |delta velocity|."""
        if trajs_vel is None:
            x = _cuda3(trajs, "compute_smoothness")
            return _path_sums(x, 2, x.shape[2])
        v = _cuda3(trajs_vel, "compute_smoothness")
        return _path_sums(v, 0, v.shape[2])

    @staticmethod
    def compute_path_length(trajectories):
        """metrics.py:33-45 — [batch] path lengths over the (x, y) columns; 0.0 for an empty batch."""
        if torch.is_tensor(trajectories) and trajectories.dim() == 3 and len(trajectories) == 0:
            return torch.tensor(0.0, device=trajectories.device)
        return _path_sums(_cuda3(trajectories, "compute_path_length"), 0, 2)

    @staticmethod
    def compute_collision_intensity(trajs, box_centers, box_sizes):
        """metrics.py:48-82 — [batch] fraction of waypoints inside any box."""
        x = _cuda3(trajs, "compute_collision_intensity")
        c = torch.as_tensor(box_centers, dtype=torch.float32, device=x.device).reshape(-1, 2).contiguous()
        s = torch.as_tensor(box_sizes, dtype=torch.float32, device=x.device).contiguous()
        n = c.shape[0]
        if s.dim() == 1:
            stride = 1
        else:
            s, stride = s.reshape(n, 2).contiguous(), 2
        if s.shape[0] != n:
            raise RampError("box_sizes must be [n_boxes] or [n_boxes, 2]")
        B, H, D = x.shape
        out = torch.empty(B, device=x.device)
        check(lib().ramp_collision_intensity(ptr(x), B, H, D, ptr(c), ptr(s), stride, n, ptr(out), stream_ptr()),
              "ramp_collision_intensity")
        return out

    def trajectory_success_and_metrics(self, trajs_final, collision_intensities, threshold=0.01):
        """metrics.py:84-126 — same keys, same None / 0.0 conventions."""
        ok = collision_intensities <= threshold
        free = trajs_final[torch.where(ok)[0]]
        n_free = len(free)
        out = {"success": 1 if n_free > 0 else 0,
               "collision_intensity": collision_intensities.mean().item() * 100,
               "path_length": None, "path_length_std": None, "waypoint_variance": None,
               "free_trajectories": free, "n_free_trajectories": n_free}
        if n_free > 0:
            lengths = self.compute_path_length(free)
            st = torch.empty(3, device=lengths.device)
            check(lib().ramp_vector_stats(ptr(lengths), n_free, ptr(st), stream_ptr()), "ramp_vector_stats")
            mean, std, _ = st.tolist()
            out["path_length"], out["path_length_std"] = mean, std
            if n_free == 1:
                out["waypoint_variance"] = 0.0
            else:
                var = float(self.compute_variance_waypoints(free))
                out["waypoint_variance"] = None if np.isnan(var) else var
        return out


class DynamicMetrics(Metrics):
    """metrics.py:128-171."""

    def calculate_single_episode_metrics(self, chain_start, chain_obs, start_state_pos, goal_state_pos,
                                         goal_safe_threshold, static_collision, pursuer_radius):
        goal = goal_state_pos.cpu().numpy() if torch.is_tensor(goal_state_pos) else goal_state_pos
        reach = pursuer_radius + 0.02                      # capture radius = pursuer radius + safety margin
        # the pursuer state of tick i is compared with the evader state two entries later (metrics.py:142-150)
        caught = any(np.linalg.norm(chain_start[i + 2] - chain_obs[i]) <= reach
                     for i in range(len(chain_obs)) if i + 2 < len(chain_start))
        captured = bool(static_collision or caught)
        goal_reached = bool(np.linalg.norm(chain_start[-1] - goal) <= goal_safe_threshold) and not captured
        length = 0
        for a, b in zip(chain_start[:-1], chain_start[1:]):
            length += np.linalg.norm(b - a)
        return {"static_collision": static_collision, "pursuer_capture": caught, "captured": captured,
                "goal_reached": goal_reached, "path_length": None if captured else length,
                "score": 0.5 * float(goal_reached) + 0.5 * float(not captured)}
