"""Experiment (round 2): does running the headline batch as TWO independent half-batches on two CUDA streams fill the
tails / prologues of the persistent GEMM launches?  (Each launch owns every SM with one ~225 KB CTA; the last wave of a
launch leaves SMs idle until the next launch of the same dependent chain starts.)
    python scripts/exp_two_streams.py            # prints ms per 25-step sampling: one stream vs two streams"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import G, H, D, T, build_model  # noqa: E402
from ramp_b200 import ops  # noqa: E402
from ramp_b200.testing import synthetic_cloud_2d  # noqa: E402

dev = torch.device("cuda:0")
groups = synthetic_cloud_2d(seed=5, groups=G).to(dev)
hc = {0: torch.tensor([-0.8, -0.8, 0.0, 0.0], device=dev), H - 1: torch.tensor([0.8, 0.8, 0.0, 0.0], device=dev)}


def sampler(model, B):
    pins = ops.Pins.from_hard_conds(model._repeat_hard_conds(hc, B), B, H, D, dev)
    chain = torch.empty(T + 1, B, H, D, device=dev)

    def start():
        chain[0] = torch.randn(B, H, D, device=dev)
        ops.inpaint(chain[0], pins)

    def step(k):
        t = T - 1 - k
        noise = torch.randn(B, H, D, device=dev)
        model._ddpm_step(chain[k], t, groups, noise if t > 0 else None, 0.5, compose=True, pins=pins, out=chain[k + 1])

    return start, step


def timeit(fn, reps=3):
    fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    parts = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    one = build_model(dev, 7168)
    s1, st1 = sampler(one, 1024)

    def single():
        s1()
        for k in range(T):
            st1(k)

    ms_one = timeit(single)
    one.model.release_engine()
    del one
    torch.cuda.empty_cache()
    Bp = 1024 // parts
    models = [build_model(dev, Bp * (G + 1)) for _ in range(parts)]
    streams = [torch.cuda.Stream() for _ in range(parts)]
    fns = [sampler(m, Bp) for m in models]

    def multi():
        cur = torch.cuda.current_stream()
        for s in streams:
            s.wait_stream(cur)
        for s, (start, _) in zip(streams, fns):
            with torch.cuda.stream(s):
                start()
        for k in range(T):
            for s, (_, step) in zip(streams, fns):
                with torch.cuda.stream(s):
                    step(k)
        for s in streams:
            cur.wait_stream(s)

    ms_multi = timeit(multi)
    print("one stream, 1024 trajectories: %.1f ms per sampling; %d streams x %d trajectories: %.1f ms (%.3fx)"
          % (ms_one, parts, Bp, ms_multi, ms_one / ms_multi))


if __name__ == "__main__":
    main()
