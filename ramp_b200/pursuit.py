"""The pursuer and the obstacle samplers of the dynamic planner, on the device.

In the reference every replan tick leaves the GPU for the environment: `MultiSphereFieldDynamics.update_centers`
(deps/torch_robotics/torch_robotics/environments/primitives.py:100-107) copies the centre to numpy, calls the pursuit
law (scripts/inference/core/utils.py:84-137), and `generate_sphere_points` (APFhelper_dynamic.py:18-39) samples the
pursuer disc with numpy before a KD-tree is rebuilt.  `DevicePursuer` keeps the centre in HBM and does both in ONE
kernel launch (`ramp_pursuer_tick`), writing straight into the static buffers of `graphs.ReplanGraph`; it also offers
the `centers` / `radii` / `update_centers` protocol of the reference field, so `DynamicGaussianDiffusionModel` can use
it as `dataset.env.obj_extra_list[0].fields[0]`.

numpy's random draws are inputs: `draw_disc_tape` / `draw_box_tape` take them from `np.random` in the reference's call
order, so a seeded run reproduces the reference's clouds (`tests/test_pursuit.py`)."""
import ctypes

import numpy as np
import torch

from . import ops
from ._lib import RampError, check, lib, ptr

_d, _i, _f = ctypes.c_double, ctypes.c_int, ctypes.c_float


def draw_disc_tape(n_ticks, num_points, surface_ratio=0.9):
    """[n_ticks, 2 * n_in] float64: per tick the draws generate_sphere_points makes (APFhelper_dynamic.py:28-29):
    n_in of U(0, 1) for the radius, then n_in of U(0, 2π) for the angle."""
    n_in = num_points - int(num_points * surface_ratio)
    rows = [np.concatenate((np.random.uniform(0, 1, n_in), np.random.uniform(0, 2 * np.pi, n_in))) for _ in range(n_ticks)]
    return np.asarray(rows, dtype=np.float64).reshape(n_ticks, 2 * n_in)


def draw_box_tape(n_boxes, num_points):
    """(n_edge [n_boxes] int32, u_edge [n_boxes, num_points], u_in [n_boxes, num_points, 2]) drawn box by box in the
    order of generate_box_points (APFhelper_dynamic.py:49,55,64): randint, rand(n_edge), rand(n_in, 2)."""
    n_edge = np.zeros(n_boxes, dtype=np.int32)
    u_edge = np.zeros((n_boxes, num_points), dtype=np.float64)
    u_in = np.zeros((n_boxes, num_points, 2), dtype=np.float64)
    for b in range(n_boxes):
        n_edge[b] = np.random.randint(2 * num_points // 3, num_points + 1)
        u_edge[b, :n_edge[b]] = np.random.rand(n_edge[b])
        u_in[b, :num_points - n_edge[b]] = np.random.rand(num_points - n_edge[b], 2)
    return n_edge, u_edge, u_in


def box_points(centers, sizes, num_points, tape=None, device=None):
    """generate_box_points for every box in one launch -> [n_boxes * num_points, 2] fp32 on the device.
    centers / sizes: [n_boxes, 2] float32 (the dynamic loop passes float32 arrays, diffusion_model_dynamic.py:516-517)."""
    c = torch.as_tensor(np.asarray(centers) if not torch.is_tensor(centers) else centers, dtype=torch.float32)
    s = torch.as_tensor(np.asarray(sizes) if not torch.is_tensor(sizes) else sizes, dtype=torch.float32)
    device = device or (c.device if c.is_cuda else torch.device("cuda", torch.cuda.current_device()))
    if torch.device(device).type != "cuda":
        raise RampError("box_points runs on a CUDA device (no CPU path)")
    nb = c.shape[0]
    n_edge, u_edge, u_in = draw_box_tape(nb, num_points) if tape is None else tape
    c, s = c.to(device).contiguous(), s.to(device).contiguous()
    ne = torch.as_tensor(n_edge, dtype=torch.int32).to(device)
    ue = torch.as_tensor(u_edge, dtype=torch.float64).to(device).contiguous()
    ui = torch.as_tensor(u_in, dtype=torch.float64).to(device).contiguous()
    out = torch.empty(nb * num_points, 2, device=device)
    check(lib().ramp_box_points(ptr(c), ptr(s), nb, num_points, ptr(ne), ptr(ue), ptr(ui), ptr(out), ops._stream()),
          "ramp_box_points")
    return out


class DevicePursuer:
    """Pursuer state in HBM: centre [2] fp32, tick counter, tape of interior draws.

    `tick(robot, points=, center_out=)` = update_centers + generate_sphere_points in one launch; graph-capturable (the
    time index is a device counter that the kernel advances, so every replay moves the pursuer one tick further).
    Law parameters as DynamicsGenerator.create_pursuit_dynamics(velocity_max, pursuit_strength, random_strength)."""

    dt = 0.1                                                    # core/utils.py:96

    def __init__(self, center, radius=0.1, num_points=64, velocity_max=0.5, pursuit_strength=0.8, random_strength=0.2,
                 tape=None, tape_ticks=256, surface_ratio=0.9, device=None):
        device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if device.type != "cuda":
            raise RampError("DevicePursuer lives on a CUDA device (no CPU path)")
        self.device = device
        self.radius = float(radius)
        self.num_points = int(num_points)
        self.n_rim = int(num_points * surface_ratio)
        self.n_in = self.num_points - self.n_rim
        v = np.array([[velocity_max / np.sqrt(2), velocity_max / np.sqrt(2)]])       # core/utils.py:89
        self.velocity = torch.as_tensor(v, dtype=torch.float32)
        self.pursuit_strength, self.random_strength = float(pursuit_strength), float(random_strength)
        self.center = torch.as_tensor(np.asarray(center, dtype=np.float64).reshape(-1)[:2], dtype=torch.float32).to(device)
        self.radii = torch.full((1,), self.radius, device=device)
        self.counter = torch.zeros(1, dtype=torch.int32, device=device)
        tape = draw_disc_tape(tape_ticks, num_points, surface_ratio) if tape is None else np.asarray(tape, dtype=np.float64)
        if tape.ndim != 2 or tape.shape[1] != 2 * self.n_in:
            raise RampError("disc tape must be [ticks, 2 * n_interior]")
        self.tape = torch.as_tensor(tape).to(device).contiguous()
        self.points = torch.empty(self.num_points, 2, device=device)

    # -- the reference field protocol (primitives.py:90-107) -------------------------------------------------------
    @property
    def centers(self):
        """[1, 2] snapshot (the reference field rebinds `centers` to a new tensor on every update, and the planner keeps the
        old ones in `chain_obs`, diffusion_model_dynamic.py:612)."""
        return self.center.reshape(1, 2).clone()

    def update_centers(self, t, current_state=None):
        """Move only (no points), at explicit time t, chasing current_state[0, :2] (a device tensor or host values)."""
        if t is None or current_state is None:
            return
        robot = torch.as_tensor(current_state, dtype=torch.float32).reshape(-1)[:2].to(self.device).contiguous()
        self._launch(robot, float(t), None, True, False, None, None)

    # -- one tick -------------------------------------------------------------------------------------------------
    def tick(self, robot, points=None, center_out=None, step=True, advance=True, t0=0.0):
        """robot: device fp32 tensor whose first two elements are the chased position.  Writes the disc cloud to
        `points` (default: self.points) and the centre to `center_out`.  Time = t0 + counter."""
        if not robot.is_cuda or robot.dtype != torch.float32:
            raise RampError("DevicePursuer.tick: robot position must be a CUDA fp32 tensor")
        points = self.points if points is None else points
        if points.shape[0] != self.num_points or not points.is_contiguous():
            raise RampError("DevicePursuer.tick: points buffer must be contiguous [num_points, 2]")
        self._launch(robot, float(t0), self.counter, step, advance, points, center_out)
        return points

    def _launch(self, robot, t0, counter, step, advance, points, center_out):
        v = self.velocity
        check(lib().ramp_pursuer_tick(ptr(self.center), ptr(robot), ptr(counter), _d(t0), int(step), int(advance),
                                      _f(float(v[0, 0])), _f(float(v[0, 1])), _d(self.pursuit_strength),
                                      _d(self.random_strength), _d(self.dt), _d(self.radius), self.n_rim, self.n_in,
                                      ptr(self.tape), int(self.tape.shape[0]), ptr(points), ptr(center_out),
                                      ops._stream()), "ramp_pursuer_tick")
