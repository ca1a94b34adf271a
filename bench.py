#!/usr/bin/env python
"""Benchmark of the RAMP reverse-diffusion planning hot path on B200.

  python bench.py --gpus N --steps K --warmup W        # this repo's CUDA path (one rank per GPU under torchrun)
  python bench.py --impl reference ...                 # the reference algorithm on the host CPU (port + unmodified reference)
  python bench.py --workload sweep --gpus N ...        # BASELINE config 5: 65,536 trajectories x 1 M-point cloud, strong-scaled

A "step" is one complete 25-step reverse-diffusion sampling of one batch of trajectories.
Headline workload at N = 1 (BASELINE.json configs[1]): Maze2D compositional energy — scores summed over 6
obstacle-condition groups + the unconditional branch (7 network rows per trajectory), batch 1024,
H = 64, D = 4, 25 DDPM steps, random-init weights, synthetic clouds.  Under torchrun every rank runs the
same per-GPU batch (weak scaling); the only collective is the all-gather of per-trajectory cost terms
followed by best-of-N selection.  The other BASELINE configs are extra sections of the same JSON line
(`replan`, `replan25`, `maze3d`, `sweep`, `field`), measured on rank 0 outside the headline's timed region.
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
SWEEP_TOTAL, SWEEP_POINTS = 65536, 1 << 20


def dtype_string():
    from ramp_b200._lib import lib
    fmt = int(lib().ramp_plane_format())
    planes = {2: "fp16 hi + fp16 residual", 0: "bf16 hi + bf16 lo"}.get(fmt, "format %d" % fmt)
    return "f32 (operands as %s planes, 3 tensor-core products each, fp32 accumulate)" % planes


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["bf16_tflops_sustained"]), float(p["hbm_gbs"]), "measured"
    except Exception:
        return 1400.0, 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING a timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index, period=0.2):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag, self.period = index, [], False, period

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self.stop_flag = True
        self.join(timeout=3)
        return self.summary()

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) > 2 + k and r[2 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def _events():
    return torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def _time_gpu(fn, reps, warm=1):
    for _ in range(warm):
        fn()
    e0, e1 = _events()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def build_model(device, max_rows, compose=True, use_apf=False, groups=G):
    from ramp_b200.diffusion import StaticGaussianDiffusionModel
    from ramp_b200.testing import synthetic_state_dict
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=H, state_dim=D, max_rows=max_rows, max_scenes=groups + 1)
    model = StaticGaussianDiffusionModel(model=net, n_diffusion_steps=T, predict_epsilon=True, compose=compose, use_apf=use_apf)
    model.load_state_dict(synthetic_state_dict(model, 0))   # random-init weights, non-zero proj_out
    model.ddim = False                                       # the 25-step DDPM p_sample_loop named by north_star
    return model.to(device).eval()


def _dist_setup():
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    return dist, world, rank, local, dev


def _make_timed(dist, world, dev):
    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        """EXACTLY `steps` calls between barrier + synchronize on both sides, CUDA events, max over ranks."""
        barrier()
        e0, e1 = _events()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            allms = [torch.zeros_like(ms) for _ in range(world)]
            dist.all_gather(allms, ms)
            per_rank = [float(m) for m in allms]
        else:
            per_rank = [float(ms)]
        return max(per_rank), per_rank

    return barrier, timed


def run_ours(args):
    from ramp_b200 import ops, sharding
    from ramp_b200._lib import lib
    from ramp_b200.sample_functions import ddpm_sample_fn
    from ramp_b200.testing import synthetic_cloud_2d
    dist, world, rank, local, dev = _dist_setup()
    B = args.batch
    model = build_model(dev, max_rows=args.max_rows)
    if args.eager:
        model.cuda_graph = False
    groups_host = synthetic_cloud_2d(seed=5, groups=G).pin_memory()                # [6, 6, 64, 2]
    hc_host = {0: torch.tensor([-0.8, -0.8, 0.0, 0.0]).pin_memory(), H - 1: torch.tensor([0.8, 0.8, 0.0, 0.0]).pin_memory()}
    groups = groups_host.to(dev)
    hc = {k: v.to(dev) for k, v in hc_host.items()}
    cost_cloud = ops.CloudIndex(groups.reshape(-1, 2))
    out_host = torch.empty(B, H, D).pin_memory()
    zeros_traj = torch.zeros(H, D, device=dev)
    lib().ramp_launch_count.restype = __import__("ctypes").c_ulonglong
    gather_ms = []

    def select(final):
        terms = ops.cost_terms(final.contiguous(), cost_cloud, 0.05)
        if world > 1 and args.time_gather:
            g0, g1 = _events()
            g0.record()
            allt = sharding.gather_terms(terms, world * B)
            g1.record()
            gather_ms.append((g0, g1))
        else:
            allt = sharding.gather_terms(terms, world * B)                  # NCCL all-gather of 12 B per trajectory
        return ops.select_best(allt)

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

    barrier, timed = _make_timed(dist, world, dev)
    W = max(args.warmup, 3)
    for _ in range(W):
        resident_step()
    sampler = ClockSampler(local)
    sampler.start()
    n0 = int(lib().ramp_launch_count())
    ms, per_rank_ms = timed(resident_step, args.steps)
    launches = int(lib().ramp_launch_count()) - n0
    e2e_step()
    gather_ms.clear()
    ms_e2e, _ = timed(e2e_step, args.steps)
    clocks = sampler.stop()
    gather_mean = None
    if gather_ms:
        gather_mean = sum(a.elapsed_time(b) for a, b in gather_ms) / len(gather_ms)
    # per-kernel-class device time of one more identical step (CUDA events around every engine launch); the events are
    # recorded by the host-side launch code, so this step runs the eager loop, not the captured graph
    graph_mode = model.cuda_graph
    model.cuda_graph = False
    model.model.profile(True)
    resident_step()
    if args.dump_launches and rank == 0:
        model.model.profile_dump(args.dump_launches)
    prof_ms, prof_n = model.model.profile(False, fetch=True)
    model.cuda_graph = graph_mode

    extras = {}
    if world == 1 and not args.no_sections:          # the other BASELINE configs: single-GPU sections of the N = 1 line only
        cfg2d = cfg_b1024_ours(dev, model, groups, hc, zeros_traj)
        extras["scene_encoder"] = scene_encoder_latency(dev, model, local)
        model.model.release_engine()
        torch.cuda.empty_cache()
        if args.replan_ticks > 0:
            extras["replan"] = replan_latency(dev, args.replan_ticks, max(3, args.replan_ticks // 10))
            extras["replan25"] = replan_latency(dev, max(10, args.replan_ticks // 3), 3, high=25, low=25)
            p1 = replan_latency(dev, max(10, args.replan_ticks // 2), 3, precision=1)
            extras["replan_single_mma_non_parity"] = {k: p1[k] for k in ("p50_ms", "p95_ms", "device_p50_ms", "ticks", "batch", "ddim_steps", "mma_per_product", "parity_mode")}
            extras["replan_single_mma_non_parity"]["note"] = "hi*hi only (one MMA per product): breaks north_star's 1e-3 waypoint tolerance, reported for context, never the headline"
        extras["field"] = field_roofline(dev, local)
        extras["maze3d"] = maze3d_section(dev, local)
        extras["sweep"] = sweep_section(dev, 8192, "per-GPU share of the 65,536-trajectory sweep at 8 GPUs")
        extras["reference_gpu_eager"] = reference_gpu_eager(dev, cfg2d)
    if rank == 0:
        peak_tf, peak_bw, how = peaks()
        rows_per_step = B * (G + 1) * T
        gemm_flops = rows_per_step * F_ROW
        achieved = gemm_flops / (prof_ms["gemm"] * 1e-3) / 1e12 if prof_ms["gemm"] > 0 else 0.0
        total_prof = sum(prof_ms.values())
        # DRAM bytes per GEMM launch from the committed ncu pass over all GEMM launches of one network call (same shapes
        # as this workload; the JSON names the commit it was taken at); with it the same kernel seen against the HBM roofline
        traffic, hbm_view = None, None
        for name in ("r2_gemm_traffic.json", "r1_gemm_traffic.json"):
            tpath = os.path.join(ROOT, "profiles", name)
            if os.path.exists(tpath) and prof_n["gemm"] > 0:
                with open(tpath) as f:
                    tj = json.load(f)
                # the ncu pass covers `rows` network rows per launch; the profiled (unsplit) launches cover B * (G + 1)
                traffic = tj["dram_bytes_per_launch"] * (B * (G + 1) / float(tj.get("rows", B * (G + 1))))
                gbps = traffic * prof_n["gemm"] / (prof_ms["gemm"] * 1e-3) / 1e9
                hbm_view = {"achieved": gbps, "peak": peak_bw, "unit": "GB/s", "frac": gbps / peak_bw, "source": name,
                            "taken_at_commit": tj.get("commit"),
                            "note": "DRAM bytes of all GEMM launches of one network call (ncu, committed) / live launch time"}
                break
        line = {
            "metric": "trajectories/sec (H=64, 25 steps)", "value": world * B * args.steps / (ms * 1e-3),
            "unit": "trajectories/s", "n_gpus": world, "steps": args.steps, "warmup": W,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": dtype_string(), "data": "synthetic",
            "config": {"workload": WORKLOAD, "trajectories_per_gpu": B, "horizon": H, "state_dim": D,
                       "diffusion_steps": T, "cond_groups": G, "network_rows_per_step": B * (G + 1),
                       "cloud": "6 groups x 6 obstacles x 64 points", "weights": "random-init (seeded)",
                       "l2": "activation working set per network call (>10 GB) exceeds L2; no flush needed",
                       "loop": "eager" if args.eager else "run_inference replays the captured 25-step loop (CUDA graph) from its third call on",
                       "sharding": "batch across ranks; all-gather of (free, length, smoothness) + select_best"},
            "e2e": {"value": world * B * args.steps / (ms_e2e * 1e-3), "unit": "trajectories/s",
                    "h2d_bytes_per_step": int(groups_host.numel() * 4 + 2 * D * 4), "d2h_bytes_per_step": int(B * H * D * 4 + 32)},
            "gpu_launches": launches,
            "clocks": clocks,
            "per_rank_ms_per_step": [m / args.steps for m in per_rank_ms],
            "allgather_ms": gather_mean,
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": achieved / peak_tf, "traffic": traffic, "peak_source": how + " bf16 sustained",
                         "hbm_view": hbm_view,
                         "kernel": "gemm_tc_kernel", "launches_per_step": prof_n["gemm"],
                         "avg_launch_us": 1e3 * prof_ms["gemm"] / max(prof_n["gemm"], 1),
                         "algorithmic_flops_per_launch": gemm_flops / max(prof_n["gemm"], 1),
                         "share_of_step": prof_ms["gemm"] / total_prof if total_prof else None,
                         "class_ms": prof_ms,
                         "note": "parity mode issues 3 f16 MMAs per product (hi*hi+hi*lo+lo*hi): ceiling = peak/3"},
        }
        if args.cpu_sample > 0 and world == 1:
            line["cpu_baseline"] = cpu_baseline(sample_B=args.cpu_sample)
        line.update(extras)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------
# Our own path on BASELINE config 1's combine at B = 1024 (CFG, 2 network rows per trajectory): the like-for-like
# partner of `reference_gpu_eager` (the reference cannot run G = 6).
# ---------------------------------------------------------------------------------------------------
def cfg_b1024_ours(dev, model, groups, hc, zeros_traj, B=1024):
    from ramp_b200.sample_functions import ddpm_sample_fn
    was, was_graph = model.compose, model.cuda_graph
    model.compose, model.cuda_graph = False, True          # capture at the first (warm-up) call, time replays
    cloud = groups[0].contiguous()

    def run():
        model.run_inference(None, hc, n_samples=B, horizon=H, return_chain=True, traj_normalized=zeros_traj,
                            obstacle_pts=cloud, sample_fn=ddpm_sample_fn, noise_std_extra_schedule_fn=lambda t: 0.5)

    ms = _time_gpu(run, 3, warm=2)
    model.compose, model.cuda_graph = was, was_graph
    return {"batch": B, "network_rows_per_step": 2 * B, "ms_per_sampling": ms, "trajectories_per_s": B / (ms * 1e-3),
            "network_rows_per_s": 2 * B * T / (ms * 1e-3)}


# ---------------------------------------------------------------------------------------------------
# "Reference eager on 1x B200" (BASELINE.md §3.4): the UNMODIFIED reference (oracle/_ref, copied by
# oracle/build_ref.py) with device='cuda', fp32 with TF32 off — what the authors time with TimerCUDA around run_inference
# (scripts/inference/inference_static.py:146-157,169).  Outside every timed region of the product.
# ---------------------------------------------------------------------------------------------------
def reference_gpu_eager(dev, ours_cfg):
    try:
        from oracle.build_ref import ref_root
        root = ref_root()
        if root is None:
            return {"unavailable": "oracle/_ref not built (python oracle/build_ref.py in the build container)"}
        os.environ["RAMP_REFERENCE_ROOT"] = root
        from oracle.ref_harness import import_reference
        m = import_reference()
        from mpd.models.diffusion_models.sample_functions import ddpm_sample_fn as ref_sample_fn
        from ramp_b200.testing import synthetic_cloud_2d, synthetic_state_dict
    except Exception as e:   # noqa: BLE001
        return {"unavailable": "reference import failed: %r" % (e,)}
    import contextlib
    import io
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    out = {"kind": "reference (unmodified, oracle/_ref)", "precision": "fp32, TF32 off (matmul and cuDNN)",
           "torch": torch.__version__}

    def make(compose):
        with contextlib.redirect_stdout(io.StringIO()):
            unet = m.TemporalUnetInference(state_dim=D, n_support_points=H, unet_input_dim=32, dim_mults=m.UNET_DIM_MULTS[1])
            model = m.StaticGaussianDiffusionModel(model=unet, variance_schedule="exponential", n_diffusion_steps=T,
                                                   predict_epsilon=True, compose=compose, use_apf=False, training=False)
        model.load_state_dict(synthetic_state_dict(model, 0))
        model.ddim = False
        for p in model.parameters():
            p.requires_grad_(False)
        return model.to(dev).eval()

    hc = {0: torch.tensor([-0.8, -0.8, 0.0, 0.0], device=dev), H - 1: torch.tensor([0.8, 0.8, 0.0, 0.0], device=dev)}
    zeros = torch.zeros(H, D, device=dev)
    cases = [("config1_cfg_b32", False, 32, synthetic_cloud_2d()), ("cfg_b1024", False, 1024, synthetic_cloud_2d(seed=5, groups=G)[0]),
             ("compose_g2_b512", True, 512, synthetic_cloud_2d(seed=5, groups=2))]
    for name, compose, B, cloud in cases:
        try:
            model = make(compose)
            cl = cloud.to(dev)

            def run():
                with contextlib.redirect_stdout(io.StringIO()):
                    model.run_inference(None, hc, n_samples=B, horizon=H, return_chain=True, traj_normalized=zeros,
                                        obstacle_pts=cl, sample_fn=ref_sample_fn, noise_std_extra_schedule_fn=lambda t: 0.5)

            ms = _time_gpu(run, 2, warm=1)
            rows = (3 if compose else 2) * B
            out[name] = {"batch": B, "network_rows_per_step": rows, "ms_per_sampling": ms,
                         "trajectories_per_s": B / (ms * 1e-3), "network_rows_per_s": rows * T / (ms * 1e-3)}
            del model
            torch.cuda.empty_cache()
        except Exception as e:   # noqa: BLE001
            out[name] = {"failed": repr(e)[:200]}
    if ours_cfg and "cfg_b1024" in out and "trajectories_per_s" in out["cfg_b1024"]:
        out["ours_cfg_b1024"] = ours_cfg
        out["speedup_same_config_cfg_b1024"] = ours_cfg["trajectories_per_s"] / out["cfg_b1024"]["trajectories_per_s"]
    return out


# ---------------------------------------------------------------------------------------------------
# FP32 FMA micro-benchmark: ramp_fp32_fma_probe (csrc/field_ops.cu) fills every SM with independent FFMA
# chains held in registers and returns the achieved TFLOP/s — the FP32 peak at the clock the chip sustains right now.
# ---------------------------------------------------------------------------------------------------
def fma_peak(dev):
    import ctypes
    from ramp_b200._lib import check, lib
    out = ctypes.c_float(0.0)
    best = 0.0
    for _ in range(3):
        check(lib().ramp_fp32_fma_probe(ctypes.byref(out), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)),
              "ramp_fp32_fma_probe")
        best = max(best, float(out.value))
    return best


# ---------------------------------------------------------------------------------------------------
# Potential-field kernel roofline (north_star (2)): the nearest-point query that feeds the APF push and the
# collision test, at the sweep's cloud size.  Algorithmic work = 3d - 1 FLOPs per (waypoint, point) pair
# (SURVEY §8d).  Denominators, all taken in THIS window: the SM clock sampled while the kernel loops for >= 2 s
# (nominal peak = SMs x 128 lanes x 2 x that clock) and the FFMA micro-benchmark run right after it.
# ---------------------------------------------------------------------------------------------------
def field_roofline(dev, gpu_index):
    from ramp_b200 import ops
    out = {}
    g = torch.Generator("cpu").manual_seed(7)
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    for dim, D_ in ((2, 4), (3, 6)):
        P, Bq = SWEEP_POINTS, 8192     # the sweep's per-GPU shard: 8192 trajectories x 64 waypoints, 1 M cloud points
        centres = torch.rand(256, dim, generator=g) * 2 - 1
        pts = (centres[torch.randint(0, 256, (P,), generator=g)] + 0.05 * torch.randn(P, dim, generator=g)).to(dev)
        x = (torch.rand(Bq, H, D_, generator=g) * 2 - 1).to(dev)
        cloud = ops.CloudIndex(pts)
        dist = torch.empty(Bq, H, device=dev)
        idx = torch.empty(Bq, H, device=dev, dtype=torch.int32)
        ms1 = _time_gpu(lambda: cloud.query(x, 0.07, dist, idx), 2, warm=2)
        reps = max(5, int(2500.0 / ms1))                       # >= 2.5 s of back-to-back launches with the sampler on
        smp = ClockSampler(gpu_index, period=0.1)
        smp.start()
        ms = _time_gpu(lambda: cloud.query(x, 0.07, dist, idx), reps, warm=0)
        clk = smp.stop()
        fma = fma_peak(dev)
        pairs = float(Bq * H) * P
        mhz = clk.get("sm_mhz") or 1965
        peak = sms * 128 * 2 * mhz * 1e6 / 1e12
        ach = pairs * (3 * dim - 1) / (ms * 1e-3) / 1e12
        hbm_bytes = 16.0 * P + 4.0 * dim * Bq * H + 8.0 * Bq * H
        out["d%d" % dim] = {"waypoints": Bq * H, "cloud_points": P, "ms": ms, "launches_timed": reps,
                            "pairs_per_s": pairs / (ms * 1e-3), "achieved_tflops": ach,
                            "fp32_peak_tflops": peak, "frac": ach / peak, "sm_mhz_during_kernel": mhz,
                            "clock_reasons": clk.get("reasons"), "clock_samples": clk.get("samples"),
                            "fma_microbench_tflops": fma, "frac_of_fma_microbench": ach / fma if fma else None,
                            "frac_of_nominal_1965mhz": ach / (sms * 128 * 2 * 1.965e9 / 1e12),
                            "algorithmic_flops_per_pair": 3 * dim - 1,
                            "hbm_algorithmic_gbs": hbm_bytes / (ms * 1e-3) / 1e9}
    out["kernel"] = "nn_query_kernel (TMA-bulk staged cloud tiles, register-blocked waypoints, exact re-scan)"
    out["note"] = "frac = achieved / (SMs x 128 x 2 x SM clock sampled WHILE this kernel ran); fma_microbench = FFMA-only kernel run right after"
    return out


# ---------------------------------------------------------------------------------------------------
# Scene encoder: ObstacleEncoderSet on the 7 conditioning clouds of the headline workload (6 groups +
# an unclustered one), i.e. what a moving conditioning cloud would put on the replan path every tick.
# Algorithmic work = 2 x MACs of every Linear / attention product (0.70 GF per scene), on the FP32 CUDA cores.
# ---------------------------------------------------------------------------------------------------
def scene_encoder_latency(dev, model, gpu_index):
    from ramp_b200.testing import synthetic_cloud_2d
    enc = model.model.scene_encoder
    pts = torch.cat([synthetic_cloud_2d(seed=5, groups=G), torch.rand(1, 6, 64, 2) * 2 - 1]).to(dev)
    S = pts.shape[0]
    ms = _time_gpu(lambda: enc(pts), 20, warm=3)
    tok, hid = 384, 64
    per_block = 2 * tok * hid * 3 * hid + 2 * 2 * tok * tok * hid + 2 * tok * hid * hid + 2 * 2 * tok * hid * 4 * hid
    flops = S * (9 * per_block + 2 * tok * 3 * hid * hid + 2 * tok * 2 * hid)
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    peak = sms * 128 * 2 * 1965 * 1e6 / 1e12
    cpu_ms = 1e3 * _cpu_scene_encoder(pts.cpu())
    return {"scenes": S, "ms": ms, "launches": 8, "algorithmic_gflop": flops / 1e9,
            "achieved_tflops": flops / (ms * 1e-3) / 1e12, "fp32_peak_tflops_nominal": peak,
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
# batch 1024, DDIM steps of the 100-step schedule (diffusion_model_dynamic.py:553-598), APF on the last step
# against a 3840-point static cloud + 256-point pursuer disc, cost/selection against the same 4096 points.
#   (a) reference tick shape: the last 5 of 10 DDIM steps  (`replan`)
#   (b) north_star's "full 25-step replan": 25 of 25 DDIM steps from q_sample at t = 96  (`replan25`)
# Per tick the host uploads the new conditioning cloud (pinned -> device), replays the graph and reads the
# selection back; latency = host wall clock between synchronised points (what TimerCUDA measures).
# ---------------------------------------------------------------------------------------------------
def replan_latency(dev, ticks, warm, batch=1024, high=10, low=5, precision=3):
    import numpy as np
    from ramp_b200.apf_dynamic import generate_box_points
    from ramp_b200.dynamic import DynamicGaussianDiffusionModel
    from ramp_b200.graphs import ReplanGraph
    from ramp_b200.testing import SyntheticPursuitDataset, pursuit_cloud_2d, synthetic_state_dict
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=H, state_dim=D, max_rows=2 * batch, max_scenes=2, precision=precision)
    m = DynamicGaussianDiffusionModel(model=net, n_diffusion_steps=100, predict_epsilon=True)
    m.load_state_dict(synthetic_state_dict(m, 0))
    m = m.to(dev).eval()
    m.ddim_num_inference_steps_high, m.ddim_num_inference_steps_low = high, low
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
                            tape_ticks=warm + ticks + 8, device=dev)
    sel_host = torch.empty(8).pin_memory()
    k = 3                                               # a tick with three executed states pinned
    s = torch.linspace(0, 1, H)[:, None]
    plan = torch.cat([-0.8 + 1.6 * s, -0.8 + 1.6 * s, torch.full((H, 2), 0.2)], dim=1)
    g = ReplanGraph(m, batch, H, k, cloud, static_pts, torch.zeros(256, 2), static_pts, stream_conditioning=True,
                    pursuer=pursuer)
    cond_host = cloud.cpu().pin_memory()                # conditioning cloud [6, 64, 2], re-encoded inside the graph every tick
    g.x_clean.copy_(plan.to(dev))
    g.history.copy_(plan[:k + 1].to(dev))
    lat_ms, dev_ms, survivors = [], [], 0
    e0, e1 = _events()
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
            survivors += int(sel_host[0] >= 0)
        if int(sel_host[0]) >= 0:
            g.x_clean.copy_(g.best)                              # next tick refines the selected plan
            g.x_clean[:k + 1] = g.history
    lat_ms.sort()
    dev_ms.sort()
    q = lambda v, f: v[min(len(v) - 1, int(f * len(v)))]  # noqa: E731
    rows = batch if (batch % 2 == 0 and m.collapse_duplicate_rows) else 2 * batch
    flops = rows * low * F_ROW
    out = {"p50_ms": q(lat_ms, 0.5), "p95_ms": q(lat_ms, 0.95), "device_p50_ms": q(dev_ms, 0.5), "ticks": ticks,
           "warmup_ticks": warm, "batch": batch, "ddim_steps": low, "of_inference_steps": high,
           "mma_per_product": precision, "parity_mode": precision == 3,
           "network_rows_per_step": rows, "reference_rows_per_step": 2 * batch,
           "ticks_with_a_collision_free_plan": survivors,
           "apf_cloud_points": 3840 + 256, "cost_cloud_points": 4096, "cuda_graph": True,
           "h2d_bytes_per_tick": int(cond_host.numel()) * 4, "d2h_bytes_per_tick": 32,
           "conditioning": "cloud uploaded and re-encoded (scene encoder kernels) inside the graph every tick",
           "pursuer": "motion law + 256-point disc computed by ramp_pursuer_tick inside the graph (no host step)",
           "tensor_floor_ms": 1e3 * precision * flops / (peaks()[0] * 1e12),
           "tensor_floor_ms_reference_rows": 1e3 * 3 * (2 * batch * low * F_ROW) / (peaks()[0] * 1e12),
           "bf16_1x_floor_ms_reference_rows": 1e3 * (2 * batch * low * F_ROW) / (peaks()[0] * 1e12),
           "note": "The reference stacks the batch twice for CFG, but with an even batch its odd-row latent mask makes both "
                   "halves the same rows (diffusion_model_dynamic.py:132-145, UnetInference.py:192-195): each distinct row "
                   "is evaluated once, results bit-identical. tensor_floor = 3 split MMAs per product over the evaluated "
                   "rows at the measured sustained bf16 peak; bf16_1x_floor_ms_reference_rows is SURVEY 8d's figure "
                   "(one MMA per product over 2B rows: 14.4 ms for 5 steps, 71.9 ms for 25)"}
    m.model.release_engine()
    del g
    torch.cuda.empty_cache()
    return out


# ---------------------------------------------------------------------------------------------------
# BASELINE config 4: Maze3D — GaussianDiffusionModel3d, D = 6, CFG w = 5.75, 25 DDPM steps, B = 1024 per GPU (8192 over
# 8 GPUs), encoder cloud [25, 64, 3]; cost terms + best-of-N against a 65,536-point cloud of the same obstacles.
# ---------------------------------------------------------------------------------------------------
def maze3d_section(dev, gpu_index, B=1024):
    from ramp_b200 import ops
    from ramp_b200.diffusion import GaussianDiffusionModel3d
    from ramp_b200.sample_functions import ddpm_sample_fn
    from ramp_b200.testing import synthetic_cloud_3d, synthetic_state_dict
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=H, state_dim=6, obstacle_3d=True, max_rows=2 * B, max_scenes=2)
    m = GaussianDiffusionModel3d(model=net, n_diffusion_steps=T, predict_epsilon=True)
    m.load_state_dict(synthetic_state_dict(m, 0))
    m = m.to(dev).eval()
    m.cuda_graph = True                                                        # capture at the first (warm-up) call, time replays
    enc_cloud = synthetic_cloud_3d().to(dev)                                   # [25, 64, 3]
    gen = torch.Generator("cpu").manual_seed(11)
    centres = synthetic_cloud_3d().mean(dim=1)                                  # the same 25 obstacles, resampled densely
    big = (centres[torch.randint(0, 25, (65536,), generator=gen)] + (torch.rand(65536, 3, generator=gen) * 2 - 1) * 0.1).to(dev)
    cost_cloud = ops.CloudIndex(big)
    hc = {0: torch.tensor([-0.8, -0.8, -0.8, 0.0, 0.0, 0.0], device=dev), H - 1: torch.tensor([0.8, 0.8, 0.8, 0.0, 0.0, 0.0], device=dev)}
    zeros = torch.zeros(H, 6, device=dev)
    sel_host = torch.empty(8).pin_memory()

    def step():
        chain = m.run_inference(None, hc, n_samples=B, horizon=H, return_chain=True, traj_normalized=zeros,
                                obstacle_pts=enc_cloud, sample_fn=ddpm_sample_fn, noise_std_extra_schedule_fn=lambda t: 0.5)
        terms = ops.cost_terms(chain[-1].contiguous(), cost_cloud, 0.05)
        sel_host.copy_(ops.select_best(terms), non_blocking=True)

    ms = _time_gpu(step, 3, warm=2)
    m.cuda_graph = False                         # the per-class events are recorded by the host-side launch code: eager step
    m.model.profile(True)
    step()
    prof_ms, prof_n = m.model.profile(False, fetch=True)
    final = m.run_inference(None, hc, n_samples=B, horizon=H, return_chain=False, traj_normalized=zeros,
                            obstacle_pts=enc_cloud, sample_fn=ddpm_sample_fn, noise_std_extra_schedule_fn=lambda t: 0.5)
    d = torch.empty(B, H, device=dev)
    i = torch.empty(B, H, device=dev, dtype=torch.int32)
    q_ms = _time_gpu(lambda: cost_cloud.query(final.contiguous(), 0.05, d, i), 20, warm=3)
    peak_tf = peaks()[0]
    flops = 2 * B * T * F_ROW
    ach = flops / (prof_ms["gemm"] * 1e-3) / 1e12 if prof_ms["gemm"] > 0 else 0.0
    out = {"workload": "maze3d_cfg5.75_b1024_h64_t25_ddpm", "trajectories_per_gpu": B, "state_dim": 6,
           "network_rows_per_step": 2 * B, "encoder_cloud": "25 obstacles x 64 points", "cost_cloud_points": 65536,
           "ms_per_sampling": ms, "trajectories_per_s": B / (ms * 1e-3),
           "roofline": {"bound": "tensor", "kernel": "gemm_tc_kernel", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s",
                        "frac": ach / peak_tf, "class_ms": prof_ms},
           "cost_query": {"waypoints": B * H, "cloud_points": 65536, "ms": q_ms,
                          "achieved_tflops": B * H * 65536.0 * 8 / (q_ms * 1e-3) / 1e12},
           "ceiling_traj_per_s_3mma": peak_tf * 1e12 / 3 / (2 * T * F_ROW),
           "note": "the reference's 3-D sampler is batch-1 only (diffusion_model_3d.py:147-152); the vectorised batch is "
                   "checked against looped singles and the oracle (tests/test_sampler_gpu.py, tests/test_loops_gpu.py)"}
    m.model.release_engine()
    torch.cuda.empty_cache()
    return out


# ---------------------------------------------------------------------------------------------------
# BASELINE config 5, loop level: CFG, 25 steps, the potential field applied from forward step 2 on — three
# query -> push -> inpaint rounds per step on x0 (diffusion_model_static.py:298-319, APFhelper.py:37-104) against a
# 1,048,576-point cloud (256 Gaussian blobs, sigma 0.05).  Through the product API: StaticGaussianDiffusionModel(
# n_diffusion_steps=25, use_apf=True) with ddim_num_inference_steps = 25 (25 DDIM steps of the 25-step schedule).
# ---------------------------------------------------------------------------------------------------
def _sweep_model(dev, max_rows):
    from ramp_b200.diffusion import StaticGaussianDiffusionModel
    from ramp_b200.testing import synthetic_state_dict
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=H, state_dim=D, max_rows=max_rows, max_scenes=2)
    m = StaticGaussianDiffusionModel(model=net, n_diffusion_steps=T, predict_epsilon=True, compose=False, use_apf=True)
    m.load_state_dict(synthetic_state_dict(m, 0))
    m.ddim_num_inference_steps = T
    return m.to(dev).eval()


def _sweep_cloud(dev):
    g = torch.Generator("cpu").manual_seed(7)
    centres = torch.rand(256, 2, generator=g) * 2 - 1
    return (centres[torch.randint(0, 256, (SWEEP_POINTS,), generator=g)] + 0.05 * torch.randn(SWEEP_POINTS, 2, generator=g)).to(dev)


class _SweepRunner:
    def __init__(self, dev, B, max_rows=8192):
        from ramp_b200 import ops
        from ramp_b200.testing import synthetic_cloud_2d
        self.m = _sweep_model(dev, min(max_rows, 2 * B))
        self.B, self.dev = B, dev
        self.enc_cloud = synthetic_cloud_2d(seed=5).to(dev)                   # [6, 64, 2]: U-Net conditioning (encoder is hard-wired)
        self.big = _sweep_cloud(dev)
        self.big_index = ops.CloudIndex(self.big)
        self.hc = {0: torch.tensor([-0.8, -0.8, 0.0, 0.0], device=dev), H - 1: torch.tensor([0.8, 0.8, 0.0, 0.0], device=dev)}
        self.zeros = torch.zeros(H, D, device=dev)
        self.m.set_apf_cloud(self.big_index)                                  # the APF / cost cloud of the sweep (P = 1 M)

    def sample(self):
        return self.m.run_inference(None, self.hc, n_samples=self.B, horizon=H, return_chain=False,
                                    traj_normalized=self.zeros, obstacle_pts=self.enc_cloud)

    def terms(self, final):
        from ramp_b200 import ops
        return ops.cost_terms(final.contiguous(), self.big_index, 0.05)


def sweep_section(dev, B, label):
    r = _SweepRunner(dev, B)
    sel_host = torch.empty(8).pin_memory()
    from ramp_b200 import ops

    def step():
        final = r.sample()
        sel_host.copy_(ops.select_best(r.terms(final)), non_blocking=True)

    ms = _time_gpu(step, 1, warm=1)
    d = torch.empty(B, H, device=dev)
    i = torch.empty(B, H, device=dev, dtype=torch.int32)
    x = r.sample().contiguous()
    q_ms = _time_gpu(lambda: r.big_index.query(x, 0.07, d, i), 3, warm=1)
    n_q = (T - 2) * 3 + 1
    r.m.model.profile(True)
    step()
    prof_ms, _ = r.m.model.profile(False, fetch=True)
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    out = {"label": label, "trajectories": B, "cloud_points": SWEEP_POINTS, "steps": T, "apf_from_step": 2,
           "apf_rounds_per_step": 3, "network_rows_per_step": 2 * B, "ms_per_sampling": ms,
           "trajectories_per_s": B / (ms * 1e-3), "nn_query_ms": q_ms, "nn_queries_per_sampling": n_q,
           "nn_query_share": n_q * q_ms / ms, "network_class_ms": prof_ms,
           "nn_query_frac_of_nominal_fp32_peak": (B * H * float(SWEEP_POINTS) * 5 / (q_ms * 1e-3)) / (sms * 128 * 2 * 1.965e9),
           "gemm_frac_of_bf16_peak": (2 * B * T * F_ROW / (prof_ms["gemm"] * 1e-3) / 1e12) / peaks()[0] if prof_ms["gemm"] else None}
    r.m.model.release_engine()
    torch.cuda.empty_cache()
    return out


def run_sweep(args):
    """--workload sweep: BASELINE config 5 as the bench's own workload, STRONG scaling: 65,536 trajectories in total,
    65,536 / N per GPU, one 1 M-point cloud replicated on every GPU; a step = one 25-step sampling of the whole sweep +
    cost terms, all-gather, best-of-N."""
    from ramp_b200 import ops, sharding
    from ramp_b200._lib import lib
    dist, world, rank, local, dev = _dist_setup()
    total = args.sweep_total
    B = total // world
    r = _SweepRunner(dev, B, max_rows=args.max_rows_sweep)
    sel_host = torch.empty(8).pin_memory()
    lib().ramp_launch_count.restype = __import__("ctypes").c_ulonglong

    def step():
        final = r.sample()
        sel_host.copy_(ops.select_best(sharding.gather_terms(r.terms(final), world * B)), non_blocking=True)

    barrier, timed = _make_timed(dist, world, dev)
    W = max(args.warmup, 1)
    for _ in range(W):
        step()
    smp = ClockSampler(local)
    smp.start()
    n0 = int(lib().ramp_launch_count())
    ms, per_rank = timed(step, args.steps)
    launches = int(lib().ramp_launch_count()) - n0
    clocks = smp.stop()
    d = torch.empty(B, H, device=dev)
    i = torch.empty(B, H, device=dev, dtype=torch.int32)
    x = r.sample().contiguous()
    q_ms = _time_gpu(lambda: r.big_index.query(x, 0.07, d, i), 3, warm=1)
    if rank == 0:
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        n_q = (T - 2) * 3 + 1
        ach = B * H * float(SWEEP_POINTS) * 5 / (q_ms * 1e-3) / 1e12
        peak = sms * 128 * 2 * (clocks.get("sm_mhz") or 1965) * 1e6 / 1e12
        line = {"metric": "trajectories/sec (H=64, 25 steps)", "value": total * args.steps / (ms * 1e-3),
                "unit": "trajectories/s", "n_gpus": world, "steps": args.steps, "warmup": W, "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": dtype_string(), "data": "synthetic",
                "config": {"workload": "sweep_65536traj_x_1Mpoints_cfg_t25_apf_from_step2", "trajectories_total": total,
                           "trajectories_per_gpu": B, "cloud_points": SWEEP_POINTS, "horizon": H, "state_dim": D,
                           "diffusion_steps": T, "apf": "3 query+push+inpaint rounds per step from forward step 2 (69 queries) + 1 cost query",
                           "l2": "1 M-point cloud (16 MB packed) streams through L2; activations exceed L2",
                           "sharding": "trajectories across ranks, cloud replicated; all-gather of cost terms + select_best"},
                "e2e": None, "gpu_launches": launches, "clocks": clocks, "per_rank_ms_per_step": [m / args.steps for m in per_rank],
                "roofline": {"bound": "fp32", "kernel": "nn_query_kernel", "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                             "frac": ach / peak, "traffic": None, "ms_per_launch": q_ms, "launches_per_step": n_q,
                             "share_of_step": n_q * q_ms / (ms / args.steps),
                             "algorithmic_flops_per_launch": B * H * float(SWEEP_POINTS) * 5}}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------
# CPU arms.  The headline workload (G = 6) cannot be run by the unmodified reference (it asserts two groups,
# diffusion_model_static.py:197), so its CPU arm is the oracle port on a bounded batch; the unmodified reference
# (oracle/_ref) is timed beside it on the configurations it CAN run: BASELINE config 1 exactly (B = 32, CFG, 25 DDPM
# steps) and its own compose (G = 2), with a one-thread figure and the CPU model string.
# ---------------------------------------------------------------------------------------------------
def _port_setup():
    from oracle import ramp_oracle as O
    from ramp_b200.testing import synthetic_cloud_2d, synthetic_state_dict
    from ramp_b200.diffusion import StaticGaussianDiffusionModel
    from ramp_b200.unet import TemporalUnetInference
    net = TemporalUnetInference(n_support_points=H, state_dim=D)
    sd = synthetic_state_dict(StaticGaussianDiffusionModel(model=net, n_diffusion_steps=T, predict_epsilon=True), 0)
    return O, sd, O.make_schedule(T), synthetic_cloud_2d


def _cpu_once(sample_B, groups=G, setup=None):
    from ramp_b200.testing import tape
    O, sd, sch, cloud_fn = setup or _port_setup()
    lat = O.scene_encoder_2d(sd, cloud_fn(seed=5, groups=groups) if groups > 1 else cloud_fn(seed=5)[None])
    hc = {0: torch.tensor([-0.8, -0.8, 0.0, 0.0]), H - 1: torch.tensor([0.8, 0.8, 0.0, 0.0])}
    noise = tape(100, [(sample_B, H, D)] * (T + 1))
    t0 = time.perf_counter()
    O.sample_ddpm(sd, sch, T, lat, [2.0] * groups, hc, noise, noise_scale=0.5)
    return time.perf_counter() - t0


def _use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1: the CPU arm uses every host core this process may run on."""
    try:
        n = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        n = os.cpu_count() or 1
    torch.set_num_threads(max(1, n))
    return torch.get_num_threads()


def _cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for ln in f:
                if ln.lower().startswith("model name"):
                    return ln.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def cpu_baseline(sample_B=2):
    threads = _use_all_host_threads()
    dt = _cpu_once(sample_B)
    return {"value": sample_B / dt, "unit": "trajectories/s", "cores": threads, "kind": "port", "cpu": _cpu_model(),
            "sample": "%d trajectories x %d network rows x %d DDPM steps (same model/config), %.1f s on %d threads"
                      % (sample_B, G + 1, T, dt, threads)}


def reference_unmodified_cpu(threads):
    """The unmodified reference on this host: config 1 exactly, its own compose (G = 2), a one-thread figure."""
    try:
        from oracle.build_ref import ref_root
        root = ref_root()
        if root is None:
            return {"unavailable": "oracle/_ref not built (python oracle/build_ref.py in the build container)"}
        os.environ["RAMP_REFERENCE_ROOT"] = root
        from oracle.ref_harness import import_reference
        m = import_reference()
        from mpd.models.diffusion_models.sample_functions import ddpm_sample_fn as ref_sample_fn
        from ramp_b200.testing import synthetic_cloud_2d, synthetic_state_dict
    except Exception as e:   # noqa: BLE001
        return {"unavailable": "reference import failed: %r" % (e,)}
    import contextlib
    import io

    def make(compose):
        with contextlib.redirect_stdout(io.StringIO()):
            unet = m.TemporalUnetInference(state_dim=D, n_support_points=H, unet_input_dim=32, dim_mults=m.UNET_DIM_MULTS[1])
            model = m.StaticGaussianDiffusionModel(model=unet, variance_schedule="exponential", n_diffusion_steps=T,
                                                   predict_epsilon=True, compose=compose, use_apf=False, training=False)
        model.load_state_dict(synthetic_state_dict(model, 0))
        model.ddim = False
        for p in model.parameters():
            p.requires_grad_(False)
        return model.eval()

    hc = {0: torch.tensor([-0.8, -0.8, 0.0, 0.0]), H - 1: torch.tensor([0.8, 0.8, 0.0, 0.0])}

    def run(model, B, cloud, reps):
        best = None
        for _ in range(reps):
            t0 = time.perf_counter()
            with contextlib.redirect_stdout(io.StringIO()):
                model.run_inference(None, hc, n_samples=B, horizon=H, return_chain=True, traj_normalized=torch.zeros(H, D),
                                    obstacle_pts=cloud, sample_fn=ref_sample_fn, noise_std_extra_schedule_fn=lambda t: 0.5)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)       # first call also warms the scene cache
        return best

    out = {"kind": "reference (unmodified, oracle/_ref)", "cpu": _cpu_model(), "os_cpu_count": os.cpu_count(), "threads": threads}
    cfg = make(False)
    dt = run(cfg, 32, synthetic_cloud_2d(), 2)
    out["config1_cfg_b32_t25"] = {"seconds": dt, "trajectories_per_s": 32 / dt, "network_rows_per_s": 64 * T / dt, "threads": threads}
    torch.set_num_threads(1)
    dt1 = run(cfg, 4, synthetic_cloud_2d(), 1)
    torch.set_num_threads(threads)
    out["config1_cfg_b4_one_thread"] = {"seconds": dt1, "trajectories_per_s": 4 / dt1, "network_rows_per_s": 8 * T / dt1, "threads": 1}
    cmp_ = make(True)
    dt2 = run(cmp_, 32, synthetic_cloud_2d(seed=5, groups=2), 1)
    out["compose_g2_b32_t25"] = {"seconds": dt2, "trajectories_per_s": 32 / dt2, "network_rows_per_s": 96 * T / dt2, "threads": threads}
    return out


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = _use_all_host_threads()
    sample_B = args.ref_batch
    setup = _port_setup()
    W = args.warmup
    for _ in range(W):
        _cpu_once(sample_B, setup=setup)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        _cpu_once(sample_B, setup=setup)
    dt = time.perf_counter() - t0
    val = sample_B * args.steps / dt
    port_cfg1 = _cpu_once(32, groups=1, setup=setup)
    unmod = reference_unmodified_cpu(threads)
    line = {"impl": "reference", "metric": "trajectories/sec (H=64, 25 steps)", "value": val, "unit": "trajectories/s",
            "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": W,
            "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "trajectories_per_step": sample_B, "horizon": H, "state_dim": D,
                       "diffusion_steps": T, "cond_groups": G, "network_rows_per_step": sample_B * (G + 1),
                       "cloud": "6 groups x 6 obstacles x 64 points", "weights": "random-init (seeded)",
                       "note": "bounded sample (B = %d per step) of the headline workload; the unmodified reference asserts "
                               "G == 2 (diffusion_model_static.py:197), so G = 6 runs as the oracle port; the unmodified "
                               "reference itself is timed under `reference_unmodified` on the configs it can run" % sample_B},
            "cpu_baseline": {"value": val, "unit": "trajectories/s", "cores": threads, "kind": "port", "cpu": _cpu_model(),
                             "sample": "%d trajectories x %d rows x %d steps per step, %d steps, %d warm-up"
                                       % (sample_B, G + 1, T, args.steps, W),
                             "network_rows_per_s": val * (G + 1) * T},
            "port_on_config1_cfg_b32": {"seconds": port_cfg1, "trajectories_per_s": 32 / port_cfg1,
                                        "network_rows_per_s": 64 * T / port_cfg1,
                                        "note": "the port on the workload the unmodified reference is timed on below: same rows/s "
                                                "means the G = 6 port figure is not a slowed-down straw man"},
            "reference_unmodified": unmod,
            "e2e": {"value": val, "unit": "trajectories/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="headline", choices=["headline", "sweep"])
    ap.add_argument("--batch", type=int, default=B_PER_GPU, help="trajectories per GPU (headline)")
    ap.add_argument("--max-rows", type=int, default=7168, help="network rows per engine call (workspace size)")
    ap.add_argument("--cpu-sample", type=int, default=2, help="trajectories in the bounded CPU sample of the headline line")
    ap.add_argument("--ref-batch", type=int, default=8, help="--impl reference: trajectories per step (G = 6 port)")
    ap.add_argument("--no-sections", action="store_true", help="headline only: skip the replan / field / maze3d / sweep / eager sections")
    ap.add_argument("--replan-ticks", type=int, default=60, help="timed replan ticks of the dynamic planner (0 = skip)")
    ap.add_argument("--time-gather", action="store_true", help="N > 1: CUDA events around the all-gather of the e2e leg")
    ap.add_argument("--sweep-total", type=int, default=SWEEP_TOTAL, help="--workload sweep: trajectories over all ranks")
    ap.add_argument("--max-rows-sweep", type=int, default=8192, help="network rows per engine call of the sweep (48 GB workspace)")
    ap.add_argument("--dump-launches", default="", help="write the per-launch CUDA-event CSV of one profiled step here")
    ap.add_argument("--eager", action="store_true", help="headline: keep run_inference on the eager loop (no CUDA-graph fast path)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    if args.workload == "sweep":
        run_sweep(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
