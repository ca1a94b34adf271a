#!/usr/bin/env python
"""Benchmark of the RAMP reverse-diffusion planning hot path on B200.

  python bench.py --gpus N --steps K --warmup W        # this repo's CUDA path (one rank per GPU under torchrun)
  python bench.py --impl reference ...                 # the reference algorithm on the host CPU (oracle port)

A "step" is one complete 25-step reverse-diffusion sampling of one batch of trajectories.
Workload at N = 1 (BASELINE.json configs[1]): Maze2D compositional energy — scores summed over 6
obstacle-condition groups + the unconditional branch (7 network rows per trajectory), batch 1024,
H = 64, D = 4, 25 DDPM steps, random-init weights, synthetic clouds.  Under torchrun every rank runs the
same per-GPU batch (weak scaling); the only collective is the all-gather of per-trajectory cost terms
followed by best-of-N selection.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, D, T, G, B_PER_GPU = 64, 4, 25, 6, 1024
F_ROW = 1.93e9            # algorithmic GEMM-like FLOPs per network row per evaluation, fwd + input-bwd
WORKLOAD = "maze2d_compose6_b1024_h64_t25_ddpm"


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["bf16_tflops_sustained"]), float(p["hbm_gbs"]), "measured"
    except Exception:
        return 1400.0, 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) > 2 + k and r[2 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def build_model(device, max_rows):
    from ramp_b200.diffusion import StaticGaussianDiffusionModel
    from ramp_b200.testing import synthetic_state_dict
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=H, state_dim=D, max_rows=max_rows, max_scenes=G + 1)
    model = StaticGaussianDiffusionModel(model=net, n_diffusion_steps=T, predict_epsilon=True, compose=True)
    model.load_state_dict(synthetic_state_dict(model, 0))   # random-init weights, non-zero proj_out
    model.ddim = False                                       # the 25-step DDPM p_sample_loop named by north_star
    return model.to(device).eval()


def run_ours(args):
    import torch.distributed as dist
    from ramp_b200 import ops, sharding
    from ramp_b200._lib import lib
    from ramp_b200.sample_functions import ddpm_sample_fn
    from ramp_b200.testing import synthetic_cloud_2d
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.batch
    model = build_model(dev, max_rows=args.max_rows)
    groups_host = synthetic_cloud_2d(seed=5, groups=G).pin_memory()                # [6, 6, 64, 2]
    hc_host = {0: torch.tensor([-0.8, -0.8, 0.0, 0.0]).pin_memory(), H - 1: torch.tensor([0.8, 0.8, 0.0, 0.0]).pin_memory()}
    groups = groups_host.to(dev)
    hc = {k: v.to(dev) for k, v in hc_host.items()}
    cost_cloud = ops.CloudIndex(groups.reshape(-1, 2))
    out_host = torch.empty(B, H, D).pin_memory()
    zeros_traj = torch.zeros(H, D, device=dev)
    lib().ramp_launch_count.restype = __import__("ctypes").c_ulonglong

    def select(final):
        terms = ops.cost_terms(final.contiguous(), cost_cloud, 0.05)
        return ops.select_best(sharding.gather_terms(terms, world * B))   # NCCL all-gather of 12 B per trajectory

    def resident_step():
        chain = model.run_inference(None, hc, n_samples=B, horizon=H, return_chain=True, traj_normalized=zeros_traj,
                                    obstacle_pts=groups, sample_fn=ddpm_sample_fn,
                                    noise_std_extra_schedule_fn=lambda t: 0.5)
        return select(chain[-1])

    def e2e_step():
        g = groups_host.to(dev, non_blocking=True)
        h = {k: v.to(dev, non_blocking=True) for k, v in hc_host.items()}
        chain = model.run_inference(None, h, n_samples=B, horizon=H, return_chain=True, traj_normalized=zeros_traj,
                                    obstacle_pts=g, sample_fn=ddpm_sample_fn, noise_std_extra_schedule_fn=lambda t: 0.5)
        sel = select(chain[-1])
        out_host.copy_(chain[-1], non_blocking=True)
        best = sel.cpu()                       # device -> host read of the step's result (synchronises)
        return best

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    for _ in range(max(args.warmup, 3)):
        resident_step()
    sampler = ClockSampler(local)
    sampler.start()
    n0 = int(lib().ramp_launch_count())
    ms = timed(resident_step, args.steps)
    launches = int(lib().ramp_launch_count()) - n0
    e2e_step()
    ms_e2e = timed(e2e_step, args.steps)
    sampler.stop_flag = True
    # per-kernel-class device time of one more identical step (CUDA events around every engine launch)
    model.model.profile(True)
    resident_step()
    if args.dump_launches and rank == 0:
        model.model.profile_dump(args.dump_launches)
    prof_ms, prof_n = model.model.profile(False, fetch=True)
    sampler.join(timeout=2)

    scene = None
    if rank == 0 and not args.no_field:
        scene = scene_encoder_latency(dev, model, sampler.summary().get("sm_mhz"))
    replan = None
    if rank == 0 and args.replan_ticks > 0:
        model.model.release_engine()
        torch.cuda.empty_cache()
        replan = replan_latency(dev, args.replan_ticks, max(3, args.replan_ticks // 10))
    field = None
    if rank == 0 and not args.no_field:
        field = field_roofline(dev, sampler.summary().get("sm_mhz"))
    if rank == 0:
        peak_tf, peak_bw, how = peaks()
        rows_per_step = B * (G + 1) * T
        gemm_flops = rows_per_step * F_ROW
        achieved = gemm_flops / (prof_ms["gemm"] * 1e-3) / 1e12 if prof_ms["gemm"] > 0 else 0.0
        total_prof = sum(prof_ms.values())
        # DRAM bytes per GEMM launch from the committed ncu pass over all GEMM launches of one network call (same shapes
        # as this workload); with it the same kernel seen against the HBM roofline
        traffic, hbm_view = None, None
        tpath = os.path.join(ROOT, "profiles", "r1_gemm_traffic.json")
        if os.path.exists(tpath) and prof_n["gemm"] > 0:
            with open(tpath) as f:
                traffic = json.load(f)["dram_bytes_per_launch"]
            gbps = traffic * prof_n["gemm"] / (prof_ms["gemm"] * 1e-3) / 1e9
            hbm_view = {"achieved": gbps, "peak": peak_bw, "unit": "GB/s", "frac": gbps / peak_bw,
                        "note": "measured DRAM bytes (ncu) / live launch time; tensor and HBM floors coincide (~29 ms per network call)"}
        line = {
            "metric": "trajectories/sec (H=64, 25 steps)", "value": world * B * args.steps / (ms * 1e-3),
            "unit": "trajectories/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 (split-bf16 x3 tensor-core products, fp32 accumulate)", "data": "synthetic",
            "config": {"workload": WORKLOAD, "trajectories_per_gpu": B, "horizon": H, "state_dim": D,
                       "diffusion_steps": T, "cond_groups": G, "network_rows_per_step": B * (G + 1),
                       "cloud": "6 groups x 6 obstacles x 64 points", "weights": "random-init (seeded)",
                       "l2": "activation working set per network call (>10 GB) exceeds L2; no flush needed",
                       "sharding": "batch across ranks; all-gather of (free, length, smoothness) + select_best"},
            "e2e": {"value": world * B * args.steps / (ms_e2e * 1e-3), "unit": "trajectories/s",
                    "h2d_bytes_per_step": int(groups_host.numel() * 4 + 2 * D * 4), "d2h_bytes_per_step": int(B * H * D * 4 + 32)},
            "gpu_launches": launches,
            "clocks": sampler.summary(),
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": achieved / peak_tf, "traffic": traffic, "peak_source": how + " bf16 sustained",
                         "hbm_view": hbm_view,
                         "kernel": "gemm_tc_kernel", "launches_per_step": prof_n["gemm"],
                         "avg_launch_us": 1e3 * prof_ms["gemm"] / max(prof_n["gemm"], 1),
                         "algorithmic_flops_per_launch": gemm_flops / max(prof_n["gemm"], 1),
                         "share_of_step": prof_ms["gemm"] / total_prof if total_prof else None,
                         "class_ms": prof_ms,
                         "note": "parity mode issues 3 bf16 MMAs per product (hi*hi+hi*lo+lo*hi): ceiling = peak/3"},
        }
        if args.cpu_sample > 0 and world == 1:
            line["cpu_baseline"] = cpu_baseline(sample_B=args.cpu_sample)
        if replan is not None:
            line["replan"] = replan
        if field is not None:
            line["field"] = field
        if scene is not None:
            line["scene_encoder"] = scene
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------
# Potential-field kernel roofline (north_star (2)): the nearest-point query that feeds the APF push and the
# collision test, at the sweep's cloud size.  Algorithmic work = 3d - 1 FLOPs per (waypoint, point) pair
# (SURVEY §8d); FP32 peak = SMs x 128 lanes x 2 x the SM clock sampled under load.
# ---------------------------------------------------------------------------------------------------
def field_roofline(dev, sm_mhz):
    from ramp_b200 import ops
    out = {}
    g = torch.Generator("cpu").manual_seed(7)
    for dim, D_ in ((2, 4), (3, 6)):
        P, Bq = 1 << 20, 8192          # the sweep's per-GPU shard: 8192 trajectories x 64 waypoints, 1 M cloud points
        centres = torch.rand(256, dim, generator=g) * 2 - 1
        pts = (centres[torch.randint(0, 256, (P,), generator=g)] + 0.05 * torch.randn(P, dim, generator=g)).to(dev)
        x = (torch.rand(Bq, H, D_, generator=g) * 2 - 1).to(dev)
        cloud = ops.CloudIndex(pts)
        dist = torch.empty(Bq, H, device=dev)
        idx = torch.empty(Bq, H, device=dev, dtype=torch.int32)
        for _ in range(2):
            cloud.query(x, 0.07, dist, idx)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        reps = 5
        e0.record()
        for _ in range(reps):
            cloud.query(x, 0.07, dist, idx)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        pairs = float(Bq * H) * P
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        peak = sms * 128 * 2 * (sm_mhz or 1965) * 1e6 / 1e12
        ach = pairs * (3 * dim - 1) / (ms * 1e-3) / 1e12
        out["d%d" % dim] = {"waypoints": Bq * H, "cloud_points": P, "ms": ms, "pairs_per_s": pairs / (ms * 1e-3),
                            "achieved_tflops": ach, "fp32_peak_tflops": peak, "frac": ach / peak,
                            "algorithmic_flops_per_pair": 3 * dim - 1}
    out["kernel"] = "nn_query_kernel (TMA-bulk staged cloud tiles, register-blocked waypoints, exact re-scan)"
    return out


# ---------------------------------------------------------------------------------------------------
# Scene encoder: ObstacleEncoderSet on the 7 conditioning clouds of the headline workload (6 groups +
# an unclustered one), i.e. what a moving conditioning cloud would put on the replan path every tick.
# Algorithmic work = 2 x MACs of every Linear / attention product (0.70 GF per scene), on the FP32 CUDA cores.
# ---------------------------------------------------------------------------------------------------
def scene_encoder_latency(dev, model, sm_mhz):
    from ramp_b200.testing import synthetic_cloud_2d
    enc = model.model.scene_encoder
    pts = torch.cat([synthetic_cloud_2d(seed=5, groups=G), torch.rand(1, 6, 64, 2) * 2 - 1]).to(dev)
    S = pts.shape[0]
    for _ in range(3):
        enc(pts)
    reps = 20
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        enc(pts)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    tok, hid = 384, 64
    per_block = 2 * tok * hid * 3 * hid + 2 * 2 * tok * tok * hid + 2 * tok * hid * hid + 2 * 2 * tok * hid * 4 * hid
    flops = S * (9 * per_block + 2 * tok * 3 * hid * hid + 2 * tok * 2 * hid)
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    peak = sms * 128 * 2 * (sm_mhz or 1965) * 1e6 / 1e12
    cpu_ms = 1e3 * _cpu_scene_encoder(pts.cpu())
    return {"scenes": S, "ms": ms, "launches": 8, "algorithmic_gflop": flops / 1e9,
            "achieved_tflops": flops / (ms * 1e-3) / 1e12, "fp32_peak_tflops": peak,
            "frac": flops / (ms * 1e-3) / 1e12 / peak, "cpu_port_ms": cpu_ms,
            "kernel": "se2d_embed / se2d_qkv / se2d_block / se2d_pool (csrc/scene_enc.cu)"}


def _cpu_scene_encoder(pts):
    from oracle import ramp_oracle as O
    from ramp_b200.testing import synthetic_state_dict
    from ramp_b200.diffusion import StaticGaussianDiffusionModel
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=H, state_dim=D)
    sd = synthetic_state_dict(StaticGaussianDiffusionModel(model=net, n_diffusion_steps=T, predict_epsilon=True), 0)
    t0 = time.perf_counter()
    O.scene_encoder_2d(sd, pts)
    return time.perf_counter() - t0


# ---------------------------------------------------------------------------------------------------
# Replan latency (BASELINE.json configs[2]): one stage-II tick of the dynamic planner as a CUDA graph —
# batch 1024, 5 DDIM steps of the 100-step schedule (diffusion_model_dynamic.py:553-598), APF on the last step
# against a 3840-point static cloud + 256-point pursuer disc, cost/selection against the same 4096 points.
# Per tick the host uploads the new pursuer cloud (pinned -> device), replays the graph and reads the
# selection back; latency = host wall clock between synchronised points (what TimerCUDA measures).
# ---------------------------------------------------------------------------------------------------
def replan_latency(dev, ticks, warm, batch=1024):
    import numpy as np
    from ramp_b200.apf_dynamic import generate_box_points
    from ramp_b200.dynamic import DynamicGaussianDiffusionModel
    from ramp_b200.graphs import ReplanGraph
    from ramp_b200.testing import SyntheticPursuitDataset, pursuit_cloud_2d, synthetic_state_dict
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=H, state_dim=D, max_rows=2 * batch, max_scenes=2)
    m = DynamicGaussianDiffusionModel(model=net, n_diffusion_steps=100, predict_epsilon=True)
    m.load_state_dict(synthetic_state_dict(m, 0))
    m = m.to(dev).eval()
    ds = SyntheticPursuitDataset()
    cloud = pursuit_cloud_2d(ds).to(dev)
    np.random.seed(0)
    boxes = ds.env.obj_fixed_list[0].fields[0]
    static_pts = torch.from_numpy(np.vstack([generate_box_points(c, s, 960) for c, s in
                                             zip(boxes.centers.numpy(), boxes.sizes.numpy())])).float()
    from ramp_b200.pursuit import DevicePursuer
    # the pursuer lives on the device: its motion law and its 256-point disc are the first node of the tick graph
    # (SURVEY 8f f4); the interior-point draws are numpy's, uploaded once as a tape
    pursuer = DevicePursuer(ds.env.obj_extra_list[0].fields[0].centers[0].numpy(), radius=0.1, num_points=256,
                            tape_ticks=warm + ticks, device=dev)
    sel_host = torch.empty(8).pin_memory()
    k = 3                                               # a tick with three executed states pinned
    s = torch.linspace(0, 1, H)[:, None]
    plan = torch.cat([-0.8 + 1.6 * s, -0.8 + 1.6 * s, torch.full((H, 2), 0.2)], dim=1)
    g = ReplanGraph(m, batch, H, k, cloud, static_pts, torch.zeros(256, 2), static_pts, stream_conditioning=True,
                    pursuer=pursuer)
    cond_host = cloud.cpu().pin_memory()                # conditioning cloud [6, 64, 2], re-encoded inside the graph every tick
    g.x_clean.copy_(plan.to(dev))
    g.history.copy_(plan[:k + 1].to(dev))
    lat_ms, dev_ms = [], []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for it in range(warm + ticks):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record()
        g.cloud.copy_(cond_host, non_blocking=True)
        g.replay()
        sel_host.copy_(g.sel, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        if it >= warm:
            lat_ms.append(1e3 * (t1 - t0))
            dev_ms.append(e0.elapsed_time(e1))
        if int(sel_host[0]) >= 0:
            g.x_clean.copy_(g.best)                              # next tick refines the selected plan
            g.x_clean[:k + 1] = g.history
    lat_ms.sort()
    dev_ms.sort()
    q = lambda v, f: v[min(len(v) - 1, int(f * len(v)))]  # noqa: E731
    rows = batch if (batch % 2 == 0 and m.collapse_duplicate_rows) else 2 * batch
    flops = rows * 5 * F_ROW
    return {"p50_ms": q(lat_ms, 0.5), "p95_ms": q(lat_ms, 0.95), "device_p50_ms": q(dev_ms, 0.5), "ticks": ticks,
            "warmup_ticks": warm, "batch": batch, "ddim_steps": 5, "network_rows_per_step": rows,
            "reference_rows_per_step": 2 * batch,
            "apf_cloud_points": 3840 + 256, "cost_cloud_points": 4096, "cuda_graph": True,
            "h2d_bytes_per_tick": int(cond_host.numel()) * 4, "d2h_bytes_per_tick": 32,
            "conditioning": "cloud uploaded and re-encoded (scene encoder kernels) inside the graph every tick",
            "pursuer": "motion law + 256-point disc computed by ramp_pursuer_tick inside the graph (no host step)",
            "tensor_floor_ms": 1e3 * 3 * flops / (peaks()[0] * 1e12),
            "note": "reference tick shape (5 DDIM steps of 100). The reference stacks the batch twice for CFG, but with an "
                    "even batch its odd-row latent mask makes both halves the same rows (diffusion_model_dynamic.py:132-145, "
                    "UnetInference.py:192-195): each distinct row is evaluated once, results bit-identical. floor = 3 "
                    "split-bf16 MMAs per product over the evaluated rows at the measured sustained bf16 peak"}


# ---------------------------------------------------------------------------------------------------
# CPU baseline: the reference algorithm restated on the host (oracle port; /root/reference does not exist
# on the GPU box).  Same model, same 7 rows per trajectory, same 25 steps — on a BOUNDED batch.
# ---------------------------------------------------------------------------------------------------
def _cpu_once(sample_B):
    from oracle import ramp_oracle as O
    from ramp_b200.testing import synthetic_cloud_2d, synthetic_state_dict, tape
    from ramp_b200.diffusion import StaticGaussianDiffusionModel
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=H, state_dim=D)
    sd = synthetic_state_dict(StaticGaussianDiffusionModel(model=net, n_diffusion_steps=T, predict_epsilon=True), 0)
    sch = O.make_schedule(T)
    lat = O.scene_encoder_2d(sd, synthetic_cloud_2d(seed=5, groups=G))
    hc = {0: torch.tensor([-0.8, -0.8, 0.0, 0.0]), H - 1: torch.tensor([0.8, 0.8, 0.0, 0.0])}
    noise = tape(100, [(sample_B, H, D)] * (T + 1))
    t0 = time.perf_counter()
    O.sample_ddpm(sd, sch, T, lat, [2.0] * G, hc, noise, noise_scale=0.5)
    return time.perf_counter() - t0


def _use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1: the CPU arm uses every host core this process may run on."""
    try:
        n = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        n = os.cpu_count() or 1
    torch.set_num_threads(max(1, n))
    return torch.get_num_threads()


def cpu_baseline(sample_B=2):
    threads = _use_all_host_threads()
    dt = _cpu_once(sample_B)
    return {"value": sample_B / dt, "unit": "trajectories/s", "cores": threads, "kind": "port",
            "sample": "%d trajectories x %d network rows x %d DDPM steps (same model/config), %.1f s on %d threads"
                      % (sample_B, G + 1, T, dt, threads)}


def run_reference(args):
    """--impl reference: the reference's CPU algorithm (oracle port) on this box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = _use_all_host_threads()
    sample_B = args.cpu_sample
    for _ in range(min(args.warmup, 1)):
        _cpu_once(sample_B)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        _cpu_once(sample_B)
    dt = time.perf_counter() - t0
    val = sample_B * args.steps / dt
    line = {"impl": "reference", "metric": "trajectories/sec (H=64, 25 steps)", "value": val, "unit": "trajectories/s",
            "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": min(args.warmup, 1),
            "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "trajectories_per_step": sample_B, "horizon": H, "state_dim": D,
                       "diffusion_steps": T, "cond_groups": G,
                       "note": "bounded sample of the same workload; reference = pure PyTorch, run as the CPU oracle port"},
            "cpu_baseline": {"value": val, "unit": "trajectories/s", "cores": threads, "kind": "port",
                             "sample": "%d trajectories per step, %d steps" % (sample_B, args.steps)},
            "e2e": {"value": val, "unit": "trajectories/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=B_PER_GPU, help="trajectories per GPU")
    ap.add_argument("--max-rows", type=int, default=7168, help="network rows per engine call (workspace size)")
    ap.add_argument("--cpu-sample", type=int, default=2, help="trajectories in the bounded CPU sample")
    ap.add_argument("--no-field", action="store_true", help="skip the potential-field kernel roofline section")
    ap.add_argument("--replan-ticks", type=int, default=60, help="timed replan ticks of the dynamic planner (0 = skip)")
    ap.add_argument("--dump-launches", default="", help="write the per-launch CUDA-event CSV of one profiled step here")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    run_ours(args)


if __name__ == "__main__":
    main()
