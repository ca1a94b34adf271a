"""Thin Python wrappers over the C ABI (include/ramp_b200.h) for the sampler-step and point-cloud
kernels.  Tensors must be CUDA fp32 / int32 / uint8 and contiguous; nothing here computes on the host."""
import ctypes

import torch

from ._lib import RampError, check, lib, ptr

_c_long, _c_int, _c_float, _vp = ctypes.c_long, ctypes.c_int, ctypes.c_float, ctypes.c_void_p


def _stream():
    return _vp(torch.cuda.current_stream().cuda_stream)


def _need_cuda(*ts):
    for t in ts:
        if t is not None and (not t.is_cuda or not t.is_contiguous()):
            raise RampError("ramp_b200 ops need contiguous CUDA tensors (no CPU path)")


class StepArgs(ctypes.Structure):
    _fields_ = [("x", _vp), ("eps", _vp), ("eps_bstride", _c_long), ("eps_gstride", _c_long),
                ("B", _c_int), ("H", _c_int), ("D", _c_int), ("G", _c_int), ("w", _c_float * 8),
                ("sqrt_recip", _c_float), ("sqrt_recipm1", _c_float),
                ("coef1", _c_float), ("coef2", _c_float), ("sigma", _c_float), ("noise_scale", _c_float), ("noise", _vp),
                ("sa_t", _c_float), ("sb_t", _c_float), ("sa_prev", _c_float), ("sdir", _c_float),
                ("pin_mask", _vp), ("pin_val", _vp), ("pin_bstride", _c_long),
                ("out", _vp), ("x0_out", _vp), ("eps_out", _vp), ("mode", _c_int), ("predict_epsilon", _c_int)]


class Pins:
    """Inpainting table: waypoints with mask != 0 are overwritten by val[b or 0, h, :]."""

    def __init__(self, mask, val):
        _need_cuda(mask, val)
        self.mask, self.val = mask, val
        self.bstride = 0 if val.shape[0] == 1 else val.shape[1] * val.shape[2]

    @staticmethod
    def from_hard_conds(hard_conds, B, H, D, device):
        """{t: [D] or [B, D]} (sample_functions.py:5-10) -> Pins."""
        mask = torch.zeros(H, dtype=torch.uint8)
        per_batch = any(v.dim() == 2 and v.shape[0] > 1 and not bool((v == v[:1]).all()) for v in hard_conds.values())
        val = torch.zeros(B if per_batch else 1, H, D, device=device)
        for t, v in hard_conds.items():
            t = t % H
            mask[t] = 1
            v = v.to(device=device, dtype=torch.float32)
            val[:, t, :] = v if (per_batch or v.dim() == 1) else v[:1]
        return Pins(mask.to(device), val.contiguous())


def step(x, eps, G, weights, out, mode, pins=None, eps_bstride=None, eps_gstride=1, sqrt_recip=0.0, sqrt_recipm1=0.0,
         coef1=0.0, coef2=0.0, sigma=0.0, noise_scale=1.0, noise=None, sa_t=0.0, sb_t=1.0, sa_prev=0.0, sdir=0.0,
         x0_out=None, eps_out=None, predict_epsilon=True):
    """One fused reverse step (mode 0 DDPM / 1 DDIM / 2 x0-only / 3 noise-only: eps may be None).
    eps rows: (b, g) -> b*eps_bstride + g*eps_gstride."""
    _need_cuda(x, eps, out, noise, x0_out, eps_out)
    B, H, D = x.shape
    a = StepArgs()
    a.x, a.eps = x.data_ptr(), (0 if eps is None else eps.data_ptr())
    a.eps_bstride, a.eps_gstride = (G + 1) if eps_bstride is None else eps_bstride, eps_gstride
    a.B, a.H, a.D, a.G = B, H, D, G
    for i, w in enumerate(weights):
        a.w[i] = float(w)
    a.sqrt_recip, a.sqrt_recipm1 = float(sqrt_recip), float(sqrt_recipm1)
    a.coef1, a.coef2, a.sigma, a.noise_scale = float(coef1), float(coef2), float(sigma), float(noise_scale)
    a.noise = 0 if noise is None else noise.data_ptr()
    a.sa_t, a.sb_t, a.sa_prev, a.sdir = float(sa_t), float(sb_t), float(sa_prev), float(sdir)
    if pins is not None:
        a.pin_mask, a.pin_val, a.pin_bstride = pins.mask.data_ptr(), pins.val.data_ptr(), pins.bstride
    a.out = 0 if out is None else out.data_ptr()
    a.x0_out = 0 if x0_out is None else x0_out.data_ptr()
    a.eps_out = 0 if eps_out is None else eps_out.data_ptr()
    a.mode, a.predict_epsilon = mode, 1 if predict_epsilon else 0
    check(lib().ramp_step(ctypes.byref(a), _stream()), "ramp_step")
    return out


def ddim_update(x, x0, out, sa_t, sb_t, sa_prev, sdir, pins=None):
    _need_cuda(x, x0, out)
    B, H, D = x.shape
    check(lib().ramp_ddim_update(ptr(x), ptr(x0), B, H, D, _c_float(sa_t), _c_float(sb_t), _c_float(sa_prev),
                                 _c_float(sdir), ptr(pins.mask if pins else None), ptr(pins.val if pins else None),
                                 _c_long(pins.bstride if pins else 0), ptr(out), _stream()), "ramp_ddim_update")
    return out


def inpaint(x, pins):
    _need_cuda(x)
    B, H, D = x.shape
    check(lib().ramp_inpaint(ptr(x), B, H, D, ptr(pins.mask), ptr(pins.val), _c_long(pins.bstride), _stream()),
          "ramp_inpaint")
    return x


def q_sample(x0, noise, out, sqrt_ac, sqrt_1mac, pins=None):
    """x0: [H, D] (broadcast) or [B, H, D]."""
    _need_cuda(x0, noise, out)
    B, H, D = noise.shape
    bs = 0 if x0.dim() == 2 else H * D
    check(lib().ramp_q_sample(ptr(x0), _c_long(bs), ptr(noise), B, H, D, _c_float(sqrt_ac), _c_float(sqrt_1mac),
                              ptr(pins.mask if pins else None), ptr(pins.val if pins else None),
                              _c_long(pins.bstride if pins else 0), ptr(out), _stream()), "ramp_q_sample")
    return out


def _morton_order(pts):
    """argsort of the points along a Z-order curve over their bounding box (device-side torch ops, once per cloud)."""
    P, dim = pts.shape
    bits = 16 if dim == 2 else 10
    lo = pts.amin(dim=0, keepdim=True)
    span = (pts.amax(dim=0, keepdim=True) - lo).clamp_min(1e-30)
    cell = ((pts - lo) / span * (2 ** bits - 1)).round().to(torch.int64).clamp_(0, 2 ** bits - 1)
    code = torch.zeros(P, dtype=torch.int64, device=pts.device)
    for b in range(bits):
        for d in range(dim):
            code |= ((cell[:, d] >> b) & 1) << (b * dim + d)
    return torch.argsort(code, stable=True).to(torch.int32)


class CloudIndex:
    """A point cloud resident in HBM in the packed layout of the nearest-point kernel.  Clouds of >= SORT_MIN points are
    also kept in Morton order for the queries (`packed_sorted`, `perm`): the kernel's exact-tie fallback re-scans the whole
    cloud when three 64-point chunks tie within the rounding bound of the scan key, which on dense clouds (the 1 M-point
    sweep) is common in arrival order and rare once near-tie points — spatial neighbours — share a chunk.  Indices returned
    by `query` always refer to the caller's order; `packed` (caller's order) is what the APF kernels fetch neighbours from."""
    SORT_MIN = 16384

    def __init__(self, points):
        pts = points.detach().reshape(-1, points.shape[-1]).to(torch.float32).contiguous()
        if not pts.is_cuda:
            raise RampError("CloudIndex needs a CUDA tensor (no CPU path)")
        self.P, self.dim = pts.shape
        self.points = pts
        lib().ramp_cloud_packed_points.restype = ctypes.c_size_t
        n = int(lib().ramp_cloud_packed_points(self.P))
        self.packed = torch.empty(n, 4, device=pts.device)
        self.perm, self.packed_sorted = None, None
        self.repack()

    def repack(self):
        """(Re-)pack after `self.points` was set or updated in place (a moving pursuer).  Small clouds: one kernel,
        graph-capturable; clouds of >= SORT_MIN points are re-sorted too (torch ops)."""
        check(lib().ramp_cloud_pack(ptr(self.points), self.P, self.dim, ptr(self.packed), _stream()), "ramp_cloud_pack")
        if self.P >= self.SORT_MIN:
            self.perm = _morton_order(self.points)
            srt = self.points.index_select(0, self.perm.long()).contiguous()
            if self.packed_sorted is None:
                self.packed_sorted = torch.empty_like(self.packed)
            check(lib().ramp_cloud_pack(ptr(srt), self.P, self.dim, ptr(self.packed_sorted), _stream()), "ramp_cloud_pack")
        return self

    def query(self, x, upper_bound, dist=None, idx=None):
        """x: [..., D] trajectories (positions = first `dim` channels).  Returns (dist, idx) shaped like x[..., 0]."""
        _need_cuda(x)
        Q = x.numel() // x.shape[-1]
        dist = torch.empty(x.shape[:-1], device=x.device) if dist is None else dist
        idx = torch.empty(x.shape[:-1], device=x.device, dtype=torch.int32) if idx is None else idx
        packed = self.packed if self.perm is None else self.packed_sorted
        check(lib().ramp_nn_query(ptr(packed), ptr(self.perm), self.P, self.dim, ptr(x), Q, x.shape[-1],
                                  _c_float(upper_bound), ptr(dist), ptr(idx), _stream()), "ramp_nn_query")
        return dist, idx


_far = {}


def far_cloud(device):
    """A one-point cloud far outside the workspace (path length / smoothness without a collision test)."""
    key = str(device)
    if key not in _far:
        _far[key] = CloudIndex(torch.full((1, 2), 1.0e6, device=device))
    return _far[key]


def apf_static(x, cloud, dist, idx, thr, strength, window, out=None):
    _need_cuda(x, dist, idx)
    B, H, D = x.shape
    out = torch.empty_like(x) if out is None else out
    check(lib().ramp_apf_static(ptr(x), B, H, D, cloud.dim, ptr(dist), ptr(idx), ptr(cloud.packed), cloud.P,
                                _c_float(thr), _c_float(strength), int(window), ptr(out), _stream()), "ramp_apf_static")
    return out


def apf_dynamic(x, cloud, dist, idx, mode, window, affected, strength, thr_force, goal=None, active=None, gate_x=None,
                gate_row=0, gate_pos=None, gate_thr=0.0):
    _need_cuda(x, dist, idx, goal, active, gate_x, gate_pos)
    B, H, D = x.shape
    check(lib().ramp_apf_dynamic(ptr(x), B, H, D, ptr(dist), ptr(idx), ptr(cloud.packed), cloud.P, int(mode), int(window),
                                 int(affected), _c_float(strength), _c_float(thr_force), ptr(goal), ptr(active),
                                 ptr(gate_x), int(gate_row), ptr(gate_pos), _c_float(gate_thr), _stream()),
          "ramp_apf_dynamic")
    return x


def smooth_segment(x, start, n, dt=0.1, max_vel=0.8):
    """sm (diffusion_model_dynamic.py:192-214) in place on waypoints start+1 .. start+n."""
    _need_cuda(x)
    B, H, D = x.shape
    check(lib().ramp_smooth_segment(ptr(x), B, H, D, int(start), int(n), _c_float(dt), _c_float(max_vel), _stream()),
          "ramp_smooth_segment")
    return x


def cost_terms(x, cloud, thr, terms=None):
    """[B, 3] = (collision_free, path_length, smoothness) per trajectory (cost.py:3-54)."""
    _need_cuda(x)
    B, H, D = x.shape
    dist, _ = cloud.query(x, thr)
    terms = torch.empty(B, 3, device=x.device) if terms is None else terms
    check(lib().ramp_cost_terms(ptr(x), B, H, D, cloud.dim, ptr(dist), _c_float(thr), ptr(terms), _stream()),
          "ramp_cost_terms")
    return terms


def select_best(terms, w_smooth=0.1, w_len=0.9, normalize=True, costs_out=None, out=None):
    """-> device tensor [8]: (best index or -1, best cost, n_free, len min, len max, smooth min, smooth max,
    rank of the best among the collision-free trajectories)."""
    _need_cuda(terms, costs_out)
    out = torch.empty(8, device=terms.device) if out is None else out
    check(lib().ramp_select_best(ptr(terms), terms.shape[0], _c_float(w_smooth), _c_float(w_len), int(bool(normalize)),
                                 ptr(out), ptr(costs_out), _stream()), "ramp_select_best")
    return out
