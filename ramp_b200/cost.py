"""Drop-in for mpd/models/diffusion_models/cost.py: collision test against a point cloud, path length,
smoothness and best-of-N selection — all in CUDA kernels (ramp_nn_query, ramp_cost_terms, ramp_select_best).
The reference materialises [B, H, n_obs, n_pts] distances (cost.py:45); here a waypoint collides iff its
nearest cloud point is closer than the threshold, which is the same predicate without the tensor."""
import torch

from . import ops

_clouds = {}


def _cloud(obstacle_points):
    key = (obstacle_points.data_ptr(), tuple(obstacle_points.shape), obstacle_points._version)
    c = _clouds.get(key)
    if c is None:
        if len(_clouds) > 16:
            _clouds.clear()
        c = _clouds[key] = (ops.CloudIndex(obstacle_points), obstacle_points)   # the key tensor stays alive with its entry
    return c[0]


def _terms(trajs, obstacle_points, collision_threshold):
    assert trajs.ndim == 3
    cloud = obstacle_points if isinstance(obstacle_points, ops.CloudIndex) else _cloud(obstacle_points)
    return ops.cost_terms(trajs.contiguous().float(), cloud, collision_threshold)


def compute_path_length(trajs):
    """cost.py:3-7."""
    return _terms(trajs, ops.far_cloud(trajs.device), 0.0)[:, 1]


def compute_smoothness(trajs):
    """cost.py:19-24."""
    return _terms(trajs, ops.far_cloud(trajs.device), 0.0)[:, 2]


def compute_collision_with_pointcloud(trajs, obstacle_points, collision_threshold=0.0, safety_margin=0.05):
    """cost.py:25-54 -> bool [B], True where any waypoint is closer than the threshold to any cloud point."""
    return _terms(trajs, obstacle_points, collision_threshold)[:, 0] == 0


def compute_trajectory_costs(trajs, obstacle_points, smoothness_weight=.1, path_length_weight=.9, collision_threshold=0.0,
                             normalize=True):
    """cost.py:56-88.  Returns (best_trajectory [H, D], best_cost, total_costs [n_free], collision_free_mask [B],
    best_index within the collision-free subset), or (None, None, None, mask, None) when every trajectory collides.
    One device->host read (the "any survivor?" branch the reference also takes at cost.py:59)."""
    terms = _terms(trajs, obstacle_points, collision_threshold)
    costs = torch.empty(trajs.shape[0], device=trajs.device)
    sel = ops.select_best(terms, smoothness_weight, path_length_weight, normalize=normalize, costs_out=costs)
    mask = terms[:, 0] != 0
    sel_host = sel.cpu()
    best = int(sel_host[0])
    if best < 0:
        print('No collision-free trajectories found!')
        return None, None, None, mask, None
    best_index = torch.tensor(int(sel_host[7]), device=trajs.device)
    return trajs[best], sel[1], costs[mask], mask, best_index
