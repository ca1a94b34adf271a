"""Generate tests/golden/pursuit.pt with the UNMODIFIED reference: generate_sphere_points / generate_box_points
(mpd/models/diffusion_models/APFhelper_dynamic.py, loaded from /root/reference by path) under seeded np.random, and
MultiSphereFieldDynamics.update_centers (deps/torch_robotics) driven by the reference's pursuit law (scripts/inference/
core/utils.py).  The numpy draws are recorded by re-seeding and drawing in the same order (ramp_b200.pursuit.draw_*), so the
CUDA kernels and the oracle can be fed the very same numbers.  Build container only; test infrastructure."""
import contextlib
import importlib.util
import io
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
REF = "/root/reference"


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, path))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    import make_golden_datasets
    from ramp_b200.pursuit import draw_box_tape, draw_disc_tape
    make_golden_datasets._stub_absent_modules()                 # matplotlib, gitpython: absent here, not on the path
    sys.path.insert(0, REF)
    A = load("mpd/models/diffusion_models/APFhelper_dynamic.py", "ref_apf_dynamic")
    U = load("scripts/inference/core/utils.py", "ref_core_utils")
    with contextlib.redirect_stdout(io.StringIO()):
        import mpd.datasets  # noqa: F401  (sets up the reference's import paths for deps/torch_robotics)
        from deps.torch_robotics.torch_robotics.environments.primitives import MultiSphereFieldDynamics
    gold = {}

    # pursuer disc: a few consecutive ticks per case (the draws continue through one numpy stream)
    discs = []
    for seed, n, radius, ticks in ((1, 64, 0.1, 4), (2, 256, np.float32(0.1), 3), (3, 10, 0.25, 2), (4, 20, 0.1, 2)):
        centres = (np.random.RandomState(seed).uniform(-0.9, 0.9, (ticks, 2))).astype(np.float32)
        np.random.seed(seed)
        out = [A.generate_sphere_points(c, np.asarray(radius) if isinstance(radius, np.float32) else radius, n) for c in centres]
        np.random.seed(seed)
        tape = draw_disc_tape(ticks, n)
        discs.append(dict(seed=seed, num_points=n, radius=float(radius), centres=centres, tape=tape, points=np.stack(out)))
    gold["discs"] = discs

    # static boxes (float32 centres / sizes, as the dynamic loop passes them)
    boxes = []
    for seed, n in ((5, 64), (6, 960), (7, 9)):
        rs = np.random.RandomState(100 + seed)
        centres = rs.uniform(-1.2, 1.2, (4, 2)).astype(np.float32)
        sizes = rs.uniform(0.1, 0.6, (4, 2)).astype(np.float32)
        np.random.seed(seed)
        out = np.vstack([A.generate_box_points(c, s, n) for c, s in zip(centres, sizes)])
        np.random.seed(seed)
        n_edge, u_edge, u_in = draw_box_tape(4, n)
        boxes.append(dict(seed=seed, num_points=n, centres=centres, sizes=sizes, n_edge=n_edge, u_edge=u_edge, u_in=u_in,
                          points=out))
    gold["boxes"] = boxes

    # the reference field object moved by the reference law, float32 tensors throughout (tensor_args of the scripts)
    tracks = []
    for vmax, ps, rs_ in ((0.5, 0.8, 0.2), (0.3, 1.0, 0.0), (0.9, 0.5, 0.5)):
        fn, vel = U.DynamicsGenerator.create_pursuit_dynamics(vmax, ps, rs_)
        f = MultiSphereFieldDynamics(np.array([[-0.35, -0.95]]), np.array([0.1]), dynamics_fn=fn, velocity=vel,
                                     tensor_args=dict(device="cpu", dtype=torch.float32))
        g = torch.Generator("cpu").manual_seed(int(vmax * 100))
        robot = torch.cumsum(0.05 * torch.randn(60, 2, generator=g), dim=0) + torch.tensor([0.6, 0.7])
        track = [f.centers.clone()]
        for t in range(60):
            f.update_centers(t, robot[t:t + 1])
            track.append(f.centers.clone())
        tracks.append(dict(args=(vmax, ps, rs_), velocity=vel, robot=robot, track=torch.cat(track)))
    gold["tracks"] = tracks

    path = os.path.join(ROOT, "tests", "golden", "pursuit.pt")
    torch.save(gold, path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
