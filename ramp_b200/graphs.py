"""CUDA-graph capture of the fixed-step loops (north_star (3): "the whole fixed-step loop is captured as a CUDA
Graph for real-time replanning latency").

Every kernel of the library is enqueued on torch's current stream through the C ABI, so a loop is captured by
running it once inside `torch.cuda.graph`: per-step scalars are kernel arguments (baked into the graph nodes), the
noise is drawn by `torch.randn` inside the capture (graph-safe Philox offsets advance on every replay), and nothing in
the captured region allocates through the library (weights, workspace and tensor maps exist after the warm-up pass).

* `DdpmLoopGraph`  — `p_sample_loop` + `ddpm_sample_fn` of the static / 3-D samplers (diffusion_model_static.py:231-256,
  diffusion_model_3d.py:184-218): T steps of (score network over B*(G+1) rows, fused posterior step).
* `ReplanGraph`    — one stage-II tick of the dynamic planner (diffusion_model_dynamic.py:553-598): q_sample, pins,
  5 DDIM steps with the point-cloud potential field on the last, smoothing, cost terms, best-of-N selection.  The moving
  pursuer enters through two small static buffers (disc points, centre) refreshed by the host before each replay, or
  by `pursuit.DevicePursuer` as the first node of the graph.
"""
import torch

from . import ops
from ._lib import RampError


class _EngineGuard:
    """A captured graph holds raw pointers into the score network's workspace arena, packed weights and scene constants.
    Rebuilding the engine (a larger batch, `prepare()`, `load_state_dict()`, `.to()`, `release_engine()`) frees them, so
    every replay first checks that the engine it was captured against is still the live one and raises otherwise; the
    scene constants are re-established before the replay if an eager call conditioned the engine on another cloud."""

    def _bind(self, model):
        unet = model.model
        self._unet = unet
        self._engine_id = (unet._engine.value if unet._engine is not None else None, unet._engine_key)
        # tensors the captured kernels read that were allocated OUTSIDE the capture (regular allocator pool): keep them alive
        self._keep = [list(model._r2s.values()), getattr(model, "_groups_cache", None)]

    def _check_engine(self):
        unet = self._unet
        now = (unet._engine.value if unet._engine is not None else None, unet._engine_key)
        if now != self._engine_id:
            raise RampError("this CUDA graph was captured against a score-network engine that has since been rebuilt or "
                            "released (batch grew, prepare() / load_state_dict() / .to() / release_engine()): its workspace "
                            "pointers are stale — capture a new graph")


def _launches():
    import ctypes
    from ._lib import lib
    lib().ramp_launch_count.restype = ctypes.c_ulonglong
    return int(lib().ramp_launch_count())


def _capture(fn, warmup=2):
    """Warm up on a side stream (lazy initialisation, autotuned launch attributes, scene cache), then capture.
    Returns (graph, fn's result, number of library kernels in the graph)."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(warmup):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    n0 = _launches()
    with torch.cuda.graph(g):
        out = fn()
    return g, out, _launches() - n0


def _count_replay(n):
    import ctypes
    from ._lib import lib
    try:
        fn = lib().ramp_launch_count_add
    except AttributeError:          # an older A/B library without the counter hook
        return
    fn.restype = None
    fn(ctypes.c_ulonglong(n))


class DdpmLoopGraph(_EngineGuard):
    """Graph of one complete DDPM sampling (x_T draw, T reverse steps, inpainting) for a fixed batch."""

    def __init__(self, model, batch, horizon, hard_conds, obstacle_pts, noise_scale=0.5, extra_steps=0, own_cloud=False):
        """noise_scale: one float, or one per step (the values `noise_std_extra_schedule_fn(i)` takes in the loop).
        own_cloud: keep a private copy of the conditioning cloud (`set_cloud` then re-conditions the graph without
        touching the caller's tensor)."""
        if not next(model.parameters()).is_cuda:
            raise RampError("graphs need a CUDA model")
        self.model = model
        dev = model.betas.device
        B, H, D = batch, horizon, model.state_dim
        T = model.n_diffusion_steps
        hc = model._repeat_hard_conds(hard_conds, B)
        self.pins = ops.Pins.from_hard_conds(hc, B, H, D, dev)
        if not model.compose and obstacle_pts.dim() == 3:
            obstacle_pts = obstacle_pts.unsqueeze(0)
        self.cloud = obstacle_pts.to(dev).float().contiguous()
        if own_cloud:
            self.cloud = self.cloud.clone()
        self._own_cloud = own_cloud
        self.chain = torch.empty(T + extra_steps + 1, B, H, D, device=dev)
        steps = [max(i, 0) for i in reversed(range(-extra_steps, T))]
        scales = list(noise_scale) if isinstance(noise_scale, (list, tuple)) else [noise_scale] * len(steps)

        def loop():
            self.chain[0] = torch.randn((B, H, D), device=dev)
            ops.inpaint(self.chain[0], self.pins)
            for k, t in enumerate(steps):
                noise = torch.randn((B, H, D), device=dev)
                model._ddpm_step(self.chain[k], t, self.cloud, noise if t > 0 else None, scales[k],
                                 compose=model.compose, pins=self.pins, out=self.chain[k + 1], forward_t=k)
            return self.chain

        self.graph, _, self.n_kernels = _capture(loop)
        self._bind(model)

    def set_cloud(self, obstacle_pts):
        """New conditioning cloud of the same shape (own_cloud graphs): copied into the graph's buffer; the next replay
        re-encodes it (the scene constants are established eagerly before every replay)."""
        if not self._own_cloud:
            raise RampError("set_cloud needs a graph created with own_cloud=True")
        src = obstacle_pts if self.cloud.dim() == obstacle_pts.dim() else obstacle_pts.unsqueeze(0)
        self.cloud.copy_(src.to(self.cloud.device).float())

    def replay(self):
        """Runs the captured loop on the current stream; returns the chain [T+1, B, H, D] (static buffer).
        Raises RampError if the engine was rebuilt since the capture; re-encodes the graph's own cloud first if an
        eager call conditioned the engine on a different one in between."""
        self._check_engine()
        self.model._set_scenes(self.model._groups_and_weights(self.cloud, self.model.compose)[0])
        self.graph.replay()
        _count_replay(self.n_kernels)
        return self.chain


class ReplanGraph(_EngineGuard):
    """Graph of one replan tick of DynamicGaussianDiffusionModel for tick index k (= number of executed states - 1).

    Static inputs (update in place, then `replay()`): `x_clean` [H, D] current plan, `history` [k+1, D] executed states,
    `disc_points` [Pd, 2] + `disc_center` [2] pursuer cloud, the cost cloud whose last Pd points are the disc, and —
    with `stream_conditioning` — `cloud`, the U-Net's conditioning cloud.  With `pursuer` (a `pursuit.DevicePursuer`) the
    first node of the graph moves the pursuer toward `history[k]` and rewrites `disc_points` / `disc_center` on the
    device, so a replay needs no host-side environment step at all.
    Outputs (static): `batch` [B, H, D], `sel` [8] (ramp_select_best), `best` [H, D].
    As in the eager tick (dynamic.py, diffusion_model_dynamic.py:592-596) the pursuer disc joins the cost cloud only while
    the current state `history[k]` is within `distance_threshold_pred` of the pursuer (decided on the device).  When no
    trajectory survives the collision test, `sel[0]` is -1 and `best` is filled with NaN — the reference then re-plans
    from scratch (:596-603): `replay(sync=True)` returns None in that case, like `compute_trajectory_costs`."""

    def __init__(self, model, batch, horizon, k, obstacle_pts, static_points, disc_points, cost_static_points,
                 collision_threshold=None, stream_conditioning=False, pursuer=None):
        """stream_conditioning: re-run the scene encoder kernels on `cloud` inside the graph: update
        the static `cloud` buffer in place before a replay and the tick is conditioned on the new scene."""
        self.model = model
        dev = model.betas.device
        B, H, D = batch, horizon, model.state_dim
        self.k = k
        self.cloud = obstacle_pts.to(dev).float().contiguous()
        if stream_conditioning:
            self.cloud = self.cloud.clone()               # a static buffer of its own: the caller writes new clouds into it
        self.x_clean = torch.zeros(H, D, device=dev)
        self.history = torch.zeros(k + 1, D, device=dev)
        self.disc_points = disc_points.to(dev).float().contiguous().clone()
        self.disc_center = torch.zeros(2, device=dev)
        Pd = self.disc_points.shape[0]
        cs = cost_static_points.to(dev).float().reshape(-1, 2)
        self.cost_points = torch.cat([cs, self.disc_points], dim=0).contiguous()
        self._cost_disc = self.cost_points[cs.shape[0]:]
        self.cost_cloud = ops.CloudIndex(self.cost_points)
        self.cost_cloud.points = self.cost_points            # pack from the live buffer on every replay
        thr = model.collision_threshold_low if collision_threshold is None else collision_threshold

        class _Field:          # the three attributes _apf_x0 reads, bound to the static buffers
            pass

        f = _Field()
        f.static_cloud = ops.CloudIndex(static_points.to(dev).float())
        f.dynamic_cloud = ops.CloudIndex(self.disc_points)
        f.dynamic_cloud.points = self.disc_points
        f.dynamic_center_dev = self.disc_center
        f.distance_threshold = model.distance_threshold_static
        f.distance_threshold_pred = model.distance_threshold_pred_apf
        self.field = f
        self.pursuer = pursuer
        if pursuer is not None and pursuer.num_points != Pd:
            raise ValueError("pursuer.num_points must equal the disc buffer size")
        mask = torch.zeros(H, dtype=torch.uint8)
        mask[:k + 1] = 1
        mask[H - 1] = 1
        self.pins_q = ops.Pins(mask.to(dev), torch.zeros(1, H, D, device=dev))
        self.pins_s = ops.Pins(mask.to(dev), torch.zeros(1, H, D, device=dev))
        self.batch = torch.empty(B, H, D, device=dev)
        self.terms = torch.empty(B, 3, device=dev)
        self.sel = torch.empty(8, device=dev)
        self.best = torch.empty(H, D, device=dev)
        self._far = torch.full_like(self.disc_points, 1.0e6)          # "no pursuer in the cost cloud"
        self._nan = torch.full((H, D), float("nan"), device=dev)
        near_thr = model.distance_threshold_pred

        def tick():
            # pins from the static inputs (device-side copies: part of the graph)
            for p, zero_v in ((self.pins_q, False), (self.pins_s, True)):
                p.val[0, :k + 1] = self.history
                p.val[0, H - 1] = self.x_clean[-1]
                if zero_v:
                    p.val[0, 0, 2:] = 0.0
            if stream_conditioning:
                model._scene_sig = None                               # encode `cloud` again on this replay
            if pursuer is not None:                                   # update_centers + generate_sphere_points, one launch
                pursuer.tick(self.history[k], points=self.disc_points, center_out=self.disc_center)
            near = (self.history[k, :2] - self.disc_center).norm() < near_thr          # :592-595
            self._cost_disc.copy_(torch.where(near, self.disc_points, self._far))
            self.field.dynamic_cloud.repack()
            self.cost_cloud.repack()
            model._tick_batch(self.x_clean, None, k, k, B, self.cloud, self.field, out=self.batch, pins_q=self.pins_q,
                              pins_s=self.pins_s)
            ops.cost_terms(self.batch, self.cost_cloud, thr, terms=self.terms)
            ops.select_best(self.terms, out=self.sel)
            idx = self.sel[0].clamp(min=0).long().reshape(1)
            self.best.copy_(torch.where(self.sel[0] >= 0, self.batch.index_select(0, idx)[0], self._nan))
            self.best[0, 2:] = 0.0                                    # diffusion_model_dynamic.py:609
            return self.best

        saved = None if pursuer is None else (pursuer.center.clone(), pursuer.counter.clone())
        self.graph, _, self.n_kernels = _capture(tick)
        if saved is not None:                                         # the warm-up passes moved the pursuer: put it back
            pursuer.center.copy_(saved[0])
            pursuer.counter.copy_(saved[1])
        self._bind(model)
        self._streams = stream_conditioning

    def replay(self, sync=False):
        """Runs the tick on the current stream and returns `best` (static buffer).  sync=True reads the selection back and
        returns None when no trajectory was collision-free (the caller then takes the replan-from-scratch path)."""
        self._check_engine()
        if not self._streams:        # scene constants are not graph nodes: make sure they are still this graph's cloud
            self.model._set_scenes(self.model._groups_and_weights(self.cloud, False)[0])
        self.graph.replay()
        _count_replay(self.n_kernels)
        if self._streams:            # the replay re-encoded `cloud` behind the host-side cache's back
            self.model._scene_sig = None
        if sync and int(self.sel[0].item()) < 0:
            return None
        return self.best
