"""DynamicGaussianDiffusionModel — drop-in for mpd/models/diffusion_models/diffusion_model_dynamic.py:24-680.

Two-stage pursuit-evasion planner: stage I draws a batch with 10 DDIM steps and keeps the cheapest
collision-free trajectory; stage II re-plans every tick — re-noise the current plan (q_sample), pin the
executed history and the goal, 5 DDIM steps with the point-cloud potential field on the last one, velocity
smoothing, best-of-N selection (:495-624).  The reference does the APF with per-trajectory host KD-tree
queries in a Python loop over the batch (:417-433); here one tick is a fixed sequence of CUDA launches
(`_tick_batch`), which is what `ramp_b200.graphs.ReplanGraph` captures as a CUDA graph.  Only the
environment protocol stays on the host, as in the reference: `dataset.env.obj_extra_list[0].fields[0]
.update_centers(t, pos)` (pursuer dynamics) and the numpy RNG that samples obstacle points."""
from copy import copy

import numpy as np
import torch

from . import ops
from ._lib import RampError
from .apf_dynamic import ObstacleField, generate_sphere_points
from .cost import compute_trajectory_costs
from .diffusion import _DiffusionBase
from .pursuit import DevicePursuer
from .sample_functions import ddpm_sample_fn


class DynamicGaussianDiffusionModel(_DiffusionBase):
    cfg_weight = 2.5                                             # :157
    # APF constants (:380-387)
    obs_radius = 0.1
    points_per_obstacle = 64
    distance_threshold_static = 0.2
    distance_threshold_pred_apf = 0.4 + 0.1
    avoidance_strength_static = 0.15
    avoidance_strength_pred = 0.15
    avoidance_window_static = 8
    avoidance_window_pred = 5
    # planner constants (:526-534)
    max_iteration = 60
    safe_threshold = 0.2
    distance_threshold_pred = 0.4
    collision_threshold_high = 0.02
    collision_threshold_low = 0.05
    replan_scratch_shape = (30, 48, 4)
    sm_window_last_step = 3                                      # :573
    sm_window_after = 2                                          # :589
    collapse_duplicate_rows = True                               # evaluate identical (trajectory, latent) rows once

    def __init__(self, model=None, variance_schedule='exponential', n_diffusion_steps=100, clip_denoised=True,
                 predict_epsilon=False, loss_type='l2', context_model=None, mask_type=None, traj_len=None, **kwargs):
        kwargs.pop('compose', None)
        kwargs.pop('use_apf', None)
        kwargs.pop('training', None)
        super().__init__(model, variance_schedule, n_diffusion_steps, clip_denoised, predict_epsilon, loss_type,
                         context_model, False, False, False, **kwargs)
        self.mask_type = mask_type
        self.traj_len = traj_len
        self.ddim = True                                          # :46
        self.ddim_num_inference_steps_high = 10                   # :47
        self.ddim_num_inference_steps_low = 5                     # :48
        self._groups_cache = None

    # ------------------------------------------------------------------ conditioning
    def _groups_and_weights(self, obstacle_pts, compose=False):
        """The loop repeats the [6, 64, 2] cloud over the batch (:520-521); every row shares it -> one group."""
        key = (obstacle_pts.data_ptr(), tuple(obstacle_pts.shape), obstacle_pts._version)
        if self._groups_cache is None or self._groups_cache[0] != key:
            g = obstacle_pts
            while g.dim() > 3:
                g = g[0]
            self._groups_cache = (key, g.unsqueeze(0).to(torch.float32).contiguous(), obstacle_pts)
        return self._groups_cache[1], [self.cfg_weight]

    def _row_scenes(self, B, G, device):
        """The reference stacks the batch twice (`x.repeat(2, ...)`, :132-145) but the network zeroes the latent of every
        ODD row of that stack (`mask[1::2] = 0`, UnetInference.py:192-195).  So the "conditional" half row b is conditional
        only for even b, and the "unconditional" half row B+b is unconditional only for odd B+b: with an even batch both
        halves coincide (eps = the row itself, conditional for even b / unconditional for odd b); with an odd batch (the
        script's n_samples=35) trajectories alternate between proper and inverted guidance.  Reproduced exactly:
        network row 2b = stack row b, row 2b+1 = stack row B+b; scene 0 = cloud latent, scene 1 = zero latent."""
        b = torch.arange(B, device=device)
        first = (b % 2).to(torch.int32)
        second = ((B + b) % 2).to(torch.int32)
        return torch.stack([first, second], dim=1).reshape(-1).contiguous()

    def _eps_rows(self, x, t, G):
        """With an even batch the two stacked halves of the reference are the SAME network row twice (see _row_scenes:
        same trajectory, same latent), so only B rows are evaluated and the step kernel reads row b for both terms of
        (1+w)·ε_c − w·ε_u (eps_gstride 0).  Rows are independent of their position in the batch, so the result is
        bit-identical to evaluating the duplicate; an odd batch (the script's n_samples=35) keeps all 2B rows."""
        B = x.shape[0]
        if B % 2 or not self.collapse_duplicate_rows:
            return super()._eps_rows(x, t, G)
        key = (B, 'even', x.device)
        if key not in self._r2s:
            self._r2s = {key: (torch.arange(B, device=x.device) % 2).to(torch.int32).contiguous()}
        self._eps_strides = (1, 0)
        return self.model.energy_grad_rows(x, t, self._r2s[key], x_repeat=1)

    def p_mean_variance(self, x, hard_conds, context, t, traj_normalized=None, obstacle_pts=None, forward_t=None):
        """:152-175 — (mean, variance, log-variance, x_recon, e_comb) in DDIM mode."""
        eps = torch.empty_like(x)
        x0 = self._x0_step(x.contiguous(), int(t[0]), obstacle_pts, eps_out=eps)
        mean, var, logvar = self.q_posterior(x_start=x0, x_t=x, t=t)
        return (mean, var, logvar, x0, eps) if self.ddim else (mean, var, logvar)

    def sm(self, s1, s2, dt=0.1, num_steps=3, max_vel=.8):
        """:192-214 — [B, num_steps, 4] constant-velocity states from s1 toward s2."""
        B = s1.shape[0]
        buf = torch.empty(B, num_steps + 1, 4, device=s1.device)
        buf[:, 0] = s1
        buf[:, num_steps] = s2
        ops.smooth_segment(buf, 0, num_steps, dt, max_vel)
        return buf[:, 1:]

    def p_sample_loop(self, *a, **k):
        raise RampError("DynamicGaussianDiffusionModel.p_sample_loop is unreachable in the reference (self.ddim is "
                        "hard-wired True, diffusion_model_dynamic.py:46, and its sample_fn omits obstacle_pts, :183)")

    # ------------------------------------------------------------------ pins
    @staticmethod
    def _history_pins(history, goal_row, H, D, device, zero_start_velocity):
        """Rows 0..len(history)-1 = executed states, row H-1 = goal (:562-566, :582-587)."""
        mask = torch.zeros(H, dtype=torch.uint8)
        n = len(history)
        mask[:n] = 1
        mask[H - 1] = 1
        val = torch.zeros(1, H, D, device=device)
        val[0, :n] = torch.cat([h.reshape(1, D) for h in history], dim=0)
        val[0, H - 1] = goal_row
        if zero_start_velocity:
            val[0, 0, 2:] = 0.0
        return ops.Pins(mask.to(device), val)

    # ------------------------------------------------------------------ APF on x0 (device only)
    def _apf_x0(self, x0, x, forward_t, field, goal_row):
        """:375-435 vectorised over the batch: static push around the closest approach, then (for trajectories whose
        current state is within reach of the pursuer) the pursuer push over the whole horizon, then re-pin the goal."""
        H = x0.shape[1]
        d, i = field.static_cloud.query(x0, field.distance_threshold)
        ops.apf_dynamic(x0, field.static_cloud, d, i, 0, self.avoidance_window_static, H, self.avoidance_strength_static,
                        field.distance_threshold)
        d, i = field.dynamic_cloud.query(x0, field.distance_threshold_pred)
        ops.apf_dynamic(x0, field.dynamic_cloud, d, i, 1, self.avoidance_window_pred, H, self.avoidance_strength_pred,
                        field.distance_threshold, goal=goal_row, gate_x=x, gate_row=forward_t,
                        gate_pos=field.dynamic_center_dev, gate_thr=self.distance_threshold_pred_apf)
        x0[:, H - 1] = goal_row
        return x0

    def _make_field(self, context, device):
        """:389-404 — created once per context (static boxes are sampled with numpy's RNG at this point)."""
        if 'obstacle_field' not in context:
            sphere = context['dataset'].env.obj_extra_list[0].fields[0]
            radius = self.obs_radius

            def dynamic_obstacle_fn(t, start_pos, replan_guide=True, best_idx=None):
                if replan_guide and best_idx is not None:
                    start_pos = start_pos[best_idx].unsqueeze(0)
                sphere.update_centers(t, start_pos)
                return sphere.centers[0].cpu().numpy(), radius

            context['obstacle_field'] = ObstacleField(context['static_obstacle_centers'], context['static_obstacle_sizes'],
                                                      dynamic_obstacle_fn, self.points_per_obstacle,
                                                      distance_threshold=self.distance_threshold_static,
                                                      distance_threshold_pred=self.distance_threshold_pred_apf,
                                                      device=device,
                                                      device_pursuer=sphere if isinstance(sphere, DevicePursuer) else None)
        return context['obstacle_field']

    # ------------------------------------------------------------------ one DDIM step (reference-shaped)
    def ddim_p_sample(self, x, hard_conds, context, t, obstacle_pts, traj_normalized=None, forward_t=None, eta=0.0,
                      use_apf=False, use_clipped_model_output=False, pins=None, out=None):
        """:338-447 (eta = 0)."""
        assert use_clipped_model_output and eta == 0.0
        ti = int(t[0]) if torch.is_tensor(t) else int(t)
        t_prev = ti - self.n_diffusion_steps // self.ddim_num_inference_steps_high
        x = x.contiguous()
        if not use_apf:
            return self._ddim_step(x, ti, t_prev, obstacle_pts, pins=pins, out=out)
        x0 = self._x0_step(x, ti, obstacle_pts)
        field = self._make_field(context, x.device)
        field.update_dynamic(forward_t, x[:, forward_t, :2], replan_guide=use_apf)
        self._apf_x0(x0, x, forward_t, field, x[0, -1].clone())
        out = torch.empty_like(x) if out is None else out
        return ops.ddim_update(x, x0, out, pins=pins, **self._ddim_scalars(ti, t_prev))

    # ------------------------------------------------------------------ the replan tick, device only
    def _tick_batch(self, x_clean, history, stepp, k, B, obstacle_pts, field, noise=None, out=None, pins_q=None,
                    pins_s=None):
        """Stage II body up to (not including) the selection (:553-591): returns the batch [B, H, D].
        `field` must already hold this tick's pursuer cloud (host protocol, see ddim_p_sample_loop)."""
        H, D = x_clean.shape
        dev = x_clean.device
        s = self._sched()
        ts = self.ddim_set_timesteps(self.ddim_num_inference_steps_high)
        low = ts[-self.ddim_num_inference_steps_low:]
        goal_row = x_clean[-1].clone()
        if pins_q is None:
            pins_q = self._history_pins(history, goal_row, H, D, dev, zero_start_velocity=False)
            pins_s = self._history_pins(history, goal_row, H, D, dev, zero_start_velocity=True)
        if noise is None:
            noise = torch.randn((B, H, D), device=dev)
        x = torch.empty(B, H, D, device=dev) if out is None else out
        nt = int(low[0])
        ops.q_sample(x_clean.contiguous(), noise, x, s['sqrt_ac'][nt], s['sqrt_1mac'][nt], pins=pins_q)
        step = self.n_diffusion_steps // self.ddim_num_inference_steps_high
        for i in low:
            i = int(i)
            if i == 0:
                ops.smooth_segment(x, stepp, self.sm_window_last_step)
                x0 = self._x0_step(x, i, obstacle_pts)
                self._apf_x0(x0, x, k, field, goal_row)
                ops.ddim_update(x, x0, x, pins=pins_s, **self._ddim_scalars(i, i - step))
            else:
                self._ddim_step(x, i, i - step, obstacle_pts, pins=pins_s, out=x)
        ops.smooth_segment(x, stepp, self.sm_window_after)
        return x

    def ddim_replan_scratch(self, shape, hard_conds, context=None, traj_normalized=None, forward_t=None,
                            obstacle_pts=None, use_apf=False, executed_history=None):
        """:461-492."""
        device = self.betas.device
        B, H, D = shape
        pins_hc = ops.Pins.from_hard_conds(hard_conds, B, H, D, device)
        x = torch.randn(shape, device=device)
        ops.inpaint(x, pins_hc)
        goal_row = x[0, -1].clone()
        pins = self._history_pins(executed_history, goal_row, H, D, device, zero_start_velocity=False)
        ops.inpaint(x, pins)
        for i in self.ddim_set_timesteps(self.ddim_num_inference_steps_high):
            t = torch.full((B,), int(i), device=device, dtype=torch.long)
            x = self.ddim_p_sample(x, hard_conds, context, t, obstacle_pts, traj_normalized=traj_normalized,
                                   forward_t=forward_t, eta=0.0, use_apf=use_apf or int(i) == 0,
                                   use_clipped_model_output=True)
            if int(i) == 0:
                use_apf = True
            ops.inpaint(x, pins_hc)
            ops.inpaint(x, pins)
        return x

    # ------------------------------------------------------------------ the two-stage loop
    def ddim_p_sample_loop(self, shape, hard_conds, context=None, return_chain=False, traj_normalized=None,
                           obstacle_pts=None, t_start_guide=float('inf'), guide=None, n_guide_steps=1, **sample_kwargs):
        """:495-624."""
        device = self.betas.device
        B, H, D = shape
        pins_hc = ops.Pins.from_hard_conds(hard_conds, B, H, D, device)
        x = torch.randn(shape, device=device)
        ops.inpaint(x, pins_hc)
        dataset = context['dataset']
        context['static_obstacle_centers'] = dataset.env.obj_fixed_list[0].fields[0].centers.cpu().numpy()[:4]
        context['static_obstacle_sizes'] = dataset.env.obj_fixed_list[0].fields[0].sizes.cpu().numpy()[:4]
        sphere = dataset.env.obj_extra_list[0].fields[0]
        cloud = obstacle_pts.to(device).float().contiguous()        # [6, 64, 2]: conditioning cloud AND cost cloud
        chain_obs = []
        chain_start = [hard_conds[0][0].unsqueeze(0)]
        chain = [] if return_chain else None
        stepp = 0
        # STAGE I: high-level plan, 10 DDIM steps without APF, best of the batch (:537-550)
        step = self.n_diffusion_steps // self.ddim_num_inference_steps_high
        for i in self.ddim_set_timesteps(self.ddim_num_inference_steps_high):
            self._ddim_step(x, int(i), int(i) - step, cloud, pins=pins_hc, out=x)
        best_traj, _, _, _, _ = compute_trajectory_costs(x, cloud, collision_threshold=self.collision_threshold_high)
        if best_traj is None:
            raise RampError("stage I found no collision-free trajectory (the reference dereferences None here, :548)")
        high_plan = best_traj.clone()
        x = best_traj.clone()
        executed_history = [x[0].clone().unsqueeze(0)]
        # STAGE II: one replan per executed state (:552-621)
        for k in range(self.max_iteration):
            field = self._make_field(context, device)
            # host protocol: the pursuer reacts to the current state (row k of every trajectory = executed_history[k])
            field.update_dynamic(k, executed_history[k][:, :2].expand(B, 2), replan_guide=True)
            batch = self._tick_batch(x, executed_history, stepp, k, B, cloud, field)
            near = np.linalg.norm(batch[0, stepp, :2].cpu().numpy() - sphere.centers[0].cpu().numpy()) < self.distance_threshold_pred
            if near:
                pursuer_pts = generate_sphere_points(sphere.centers[0].cpu().numpy(), sphere.radii[0].cpu().numpy(), 64)
                cost_cloud = torch.vstack([cloud, torch.from_numpy(pursuer_pts).unsqueeze(0).to(device=device, dtype=cloud.dtype)])
            else:
                cost_cloud = cloud
            x, _, _, _, _ = compute_trajectory_costs(batch, cost_cloud, collision_threshold=self.collision_threshold_low)
            while x is None:
                print('replanning from scratch')
                n0 = self.replan_scratch_shape[0]
                new_hc = {kk: v[:n0].clone() for kk, v in hard_conds.items()}
                xs = self.ddim_replan_scratch((n0, H, D), new_hc, context, traj_normalized, forward_t=k, obstacle_pts=cloud,
                                              use_apf=False, executed_history=executed_history)
                ops.smooth_segment(xs, stepp, self.sm_window_after)
                x, _, _, _, _ = compute_trajectory_costs(xs, cloud, collision_threshold=self.collision_threshold_low)
            x = x.clone()
            x[0, 2:] = 0.0
            executed_history.append(x[stepp + 1].clone().unsqueeze(0))
            updated_start_state = x[stepp].clone()
            stepp += 1
            if return_chain:
                if stepp == 1:
                    chain.append(high_plan.unsqueeze(0).clone())
                chain.append(x.unsqueeze(0).clone())
            chain_obs.append(sphere.centers)
            chain_start.append(updated_start_state.unsqueeze(0).clone())
            if torch.norm(x[stepp - 1, :2] - x[-1, :2]) < self.safe_threshold:
                break
        if return_chain:
            chain = torch.stack(chain, dim=1)
        return x, chain, chain_obs, chain_start

    @torch.no_grad()
    def conditional_sample(self, hard_conds, horizon=None, batch_size=1, ddim=False, traj_normalized=None,
                           obstacle_pts=None, **sample_kwargs):
        """:626-636."""
        horizon = horizon or self.model.n_support_points
        for kk in ('sample_fn', 'noise_std_extra_schedule_fn', 'n_diffusion_steps_without_noise'):
            sample_kwargs.pop(kk, None)
        return self.ddim_p_sample_loop((batch_size, horizon, self.state_dim), hard_conds, traj_normalized=traj_normalized,
                                       obstacle_pts=obstacle_pts, **sample_kwargs)

    @torch.no_grad()
    def warmup(self, horizon=64, device='cuda'):
        """:642-647 — builds the engine (the reference's call passes no cloud and cannot execute)."""
        self.model.prepare(horizon=horizon)

    @torch.no_grad()
    def run_inference(self, context=None, hard_conds=None, n_samples=1, return_chain=False, traj_normalized=None,
                      obstacle_pts=None, **diffusion_kwargs):
        """:649-667 — (chain [ticks+1, 1, H, D], chain_obs, chain_start), or the last chain entry."""
        hard_conds = self._repeat_hard_conds(hard_conds, n_samples)
        context = copy(context)
        _, chain, chain_obs, chain_start = self.conditional_sample(hard_conds, context=context, batch_size=n_samples,
                                                                   return_chain=True, traj_normalized=traj_normalized,
                                                                   obstacle_pts=obstacle_pts, **diffusion_kwargs)
        chain = chain.transpose(0, 1)
        if return_chain:
            return chain, chain_obs, chain_start
        return chain[-1]
