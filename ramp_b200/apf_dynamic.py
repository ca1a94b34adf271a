"""Drop-in for mpd/models/diffusion_models/APFhelper_dynamic.py (pursuit scenes): static boxes sampled to
points + a moving pursuer disc.  Point generation stays on the host with numpy's global RNG, drawing in the
reference's order so a seeded run produces the same clouds (APFhelper_dynamic.py:18-68); the clouds are then
packed into HBM and every query / push is a CUDA kernel (ramp_nn_query, ramp_apf_dynamic)."""
import numpy as np
import torch

from . import ops

GOLDEN_ANGLE = np.pi * (3 - np.sqrt(5))


def generate_sphere_points(center, radius, num_points, surface_ratio=0.9):
    """APFhelper_dynamic.py:18-39: golden-angle rim points, then uniformly random interior points."""
    n_rim = int(num_points * surface_ratio)
    n_in = num_points - n_rim
    ang = GOLDEN_ANGLE * np.arange(n_rim)
    xs, ys = radius * np.cos(ang) + center[0], radius * np.sin(ang) + center[1]
    if n_in > 0:
        r = radius * np.sqrt(np.random.uniform(0, 1, n_in))
        th = np.random.uniform(0, 2 * np.pi, n_in)
        xs = np.concatenate((xs, r * np.cos(th) + center[0]))
        ys = np.concatenate((ys, r * np.sin(th) + center[1]))
    return np.column_stack((xs, ys))


def generate_box_points(center, size, num_points):
    """APFhelper_dynamic.py:41-68: between 2/3 and all of the points on the perimeter, the rest inside."""
    cx, cy = center
    w, h = size
    left, right, top, bottom = cx - w / 2, cx + w / 2, cy + h / 2, cy - h / 2
    n_edge = np.random.randint(2 * num_points // 3, num_points + 1)
    n_in = num_points - n_edge
    corners = np.array([[left, top], [right, top], [right, bottom], [left, bottom]])
    seg_len = np.array([w, h, w, h]).repeat(2)           # the reference walks an 8-entry length table
    s = np.random.rand(n_edge) * seg_len.sum()
    cum = np.cumsum(seg_len)
    seg = np.searchsorted(cum, s)
    frac = (s - np.concatenate(([0], cum[:-1]))[seg]) / seg_len[seg]
    a, b = corners[seg % 4], corners[(seg + 1) % 4]
    edge_pts = a + frac[:, None] * (b - a)
    inner = np.random.rand(n_in, 2)
    inner[:, 0] = inner[:, 0] * w + left
    inner[:, 1] = inner[:, 1] * h + bottom
    return np.concatenate([edge_pts, inner], axis=0)


class ObstacleField:
    """APFhelper_dynamic.py:70-105."""

    def __init__(self, static_obstacle_centers, static_obstacle_sizes, dynamic_obstacle_fn, points_per_obstacle=32,
                 distance_threshold=0.1, distance_threshold_pred=0.2, device=None, device_pursuer=None):
        self.static_obstacle_centers = static_obstacle_centers
        self.static_obstacle_sizes = static_obstacle_sizes
        self.dynamic_obstacle_fn = dynamic_obstacle_fn
        self.points_per_obstacle = points_per_obstacle
        self.distance_threshold = distance_threshold
        self.distance_threshold_pred = distance_threshold_pred
        self.last_update_time = None
        self.device_pursuer = device_pursuer
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.static_obstacle_points = np.vstack([generate_box_points(c, s, points_per_obstacle)
                                                 for c, s in zip(static_obstacle_centers, static_obstacle_sizes)])
        self.static_cloud = ops.CloudIndex(torch.as_tensor(self.static_obstacle_points, dtype=torch.float32).to(self.device))
        self.dynamic_cloud = None
        self.dynamic_points = None
        self.dynamic_center = None
        self.dynamic_center_dev = None

    def set_dynamic_points(self, points, center):
        """Install a pursuer cloud directly ([P, 2] array / tensor) — used by the graph-captured replan tick."""
        pts = torch.as_tensor(np.asarray(points), dtype=torch.float32) if not isinstance(points, torch.Tensor) else points
        self.dynamic_points = np.asarray(points) if not isinstance(points, torch.Tensor) else None
        self.dynamic_cloud = ops.CloudIndex(pts.to(self.device))
        self.dynamic_center = np.asarray(center, dtype=np.float64)
        self.dynamic_center_dev = torch.as_tensor(self.dynamic_center, dtype=torch.float32).to(self.device)

    def update_dynamic(self, t, start_pos, replan_guide=False, best_idx=None):
        if self.last_update_time != t and self.device_pursuer is not None:
            # pursuit.DevicePursuer: motion law and disc sampling are kernels, the cloud never visits the host
            p = self.device_pursuer
            if replan_guide and best_idx is not None:
                start_pos = start_pos[best_idx].unsqueeze(0)
            p.update_centers(t, start_pos)
            pts = p.tick(p.center, step=False)
            if self.dynamic_cloud is None or self.dynamic_cloud.points is not pts:
                self.dynamic_cloud = ops.CloudIndex(pts)
                self.dynamic_cloud.points = pts
            else:
                self.dynamic_cloud.repack()
            self.dynamic_points, self.dynamic_center, self.dynamic_center_dev = None, None, p.center
            self.last_update_time = t
        elif self.last_update_time != t:
            center, radius = self.dynamic_obstacle_fn(t, start_pos, replan_guide, best_idx)
            self.set_dynamic_points(generate_sphere_points(center, radius, self.points_per_obstacle), center)
            self.last_update_time = t

    def query_static(self, points):
        return self.static_cloud.query(_dev(points, self.device), self.distance_threshold)

    def query_dynamic(self, points):
        pts = _dev(points, self.device)
        if self.dynamic_cloud is None:
            n = pts.shape[0]
            return torch.full((n,), float("inf"), device=self.device), torch.arange(n, device=self.device, dtype=torch.int32)
        return self.dynamic_cloud.query(pts, self.distance_threshold_pred)


def _dev(points, device):
    pts = points if isinstance(points, torch.Tensor) else torch.as_tensor(np.asarray(points))
    return pts.to(device).float().contiguous()


def avoidance(trajectory, obstacle_field, is_dynamic=False, avoidance_window=5, avoidance_strength=0.1,
              avoidance_strength_pred=0.3, affected_states=5, goal_state=None, stepp=None):
    """APFhelper_dynamic.py:107-142 — in place on the positions of one trajectory [H, D] (or a batch [B, H, D])."""
    x = trajectory if trajectory.dim() == 3 else trajectory.unsqueeze(0)
    if not x.is_contiguous():
        raise ops.RampError("avoidance needs a contiguous trajectory (it is modified in place)")
    H = x.shape[1]
    if affected_states is None:
        affected_states = H
    if is_dynamic:
        cloud = obstacle_field.dynamic_cloud
        dist, idx = cloud.query(x, obstacle_field.distance_threshold_pred)
        ops.apf_dynamic(x, cloud, dist, idx, 1, avoidance_window, affected_states, avoidance_strength_pred,
                        obstacle_field.distance_threshold, goal=None if goal_state is None else goal_state.float().contiguous())
    else:
        cloud = obstacle_field.static_cloud
        dist, idx = cloud.query(x, obstacle_field.distance_threshold)
        ops.apf_dynamic(x, cloud, dist, idx, 0, avoidance_window, affected_states, avoidance_strength,
                        obstacle_field.distance_threshold)
    return trajectory
