"""Drop-in for mpd/models/diffusion_models/APFhelper.py: `ObstacleField` + `avoidance` (static scenes).

The reference builds a scipy cKDTree on the host and pushes colliding waypoints in a Python loop
(APFhelper.py:5-104).  Here the cloud lives in HBM in the packed layout of the nearest-point kernel and the
query + windowed Gaussian push are two CUDA launches; nothing crosses to the host."""
import numpy as np
import torch

from . import ops


class ObstacleField:
    """APFhelper.py:5-35.  obstacle_pts: [P, d] tensor or array (d = 2 or 3)."""

    def __init__(self, obstacle_pts, distance_threshold=0.1, max_leaf_size=16, device=None):
        self.distance_threshold = distance_threshold
        pts = obstacle_pts if isinstance(obstacle_pts, torch.Tensor) else torch.as_tensor(np.asarray(obstacle_pts))
        if not pts.is_cuda:
            pts = pts.to(device if device is not None else torch.device("cuda", torch.cuda.current_device()))
        pts = pts.reshape(-1, pts.shape[-1]).float().contiguous()
        self.static_obstacle_points = pts
        self.obstacle_points_tensor = pts
        self.cloud = ops.CloudIndex(pts)

    def query_static(self, points, chunk_size=1024):
        """points [Q, d] (device tensor) -> (distances [Q] (inf beyond the threshold), indices [Q] (P when none))."""
        pts = points if isinstance(points, torch.Tensor) else torch.as_tensor(np.asarray(points))
        pts = pts.to(self.cloud.points.device).float().contiguous()
        return self.cloud.query(pts, self.distance_threshold)


def avoidance(trajectories, obstacle_field, avoidance_window=7, avoidance_strength=0.2):
    """APFhelper.py:37-104 — returns a new tensor; velocities untouched."""
    x = trajectories.contiguous().float()
    dist, idx = obstacle_field.cloud.query(x, obstacle_field.distance_threshold)
    return ops.apf_static(x, obstacle_field.cloud, dist, idx, obstacle_field.distance_threshold, avoidance_strength,
                          avoidance_window)
